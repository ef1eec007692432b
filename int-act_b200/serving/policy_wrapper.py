"""`LeRobotPolicyWrapper` for the B200 engine (row a3): the object the policy server calls.

Contract of the reference wrapper (/root/reference/src/experiments/policies/policy_wrapper.py:18-170):
  * `LeRobotPolicyWrapper(pipeline_cfg=..., model_class=...)` starts with NO model and NO adapter (`:58-59`); the first
    `{"new_model_path": ...}` request calls `switch_model`, which loads the checkpoint with
    `model_class.from_pretrained(pretrained_name_or_path=path, config=model_cfg)`, moves it to the run's dtype / device,
    nulls the LM heads when `freeze_lm_head` and builds the adapter named by `eval_cfg.env_adapter_path` (`:61-67,86-113,
    141-170`);
  * `select_action(obs) -> np.float64[action_step, 7]` (or `[B, action_step, 7]` for a rollout batch) =
    env_adapter.preprocess -> `action_step` calls of `model.select_action` (queue pops, `:131-135`) stacked on dim 1 ->
    drop the batch dim when B == 1 -> env_adapter.postprocess; `reset()` resets the action queue and the adapter.
The torch.autocast / `.to(device)` of the reference's select_action have no counterpart: the adapters produce device
tensors and the engine's dtype is fixed when the model is moved with `.to(dtype)`.

Added for the B200 server: the direct constructor `LeRobotPolicyWrapper(model, env_adapter, ...)` (tests, bench) and
`select_action_many(list of obs)`, which runs several single-observation requests as ONE engine batch (pre / post-
processing stay per request, so every client gets exactly the adapter output it would have got alone).
"""
from __future__ import annotations

import gc
import threading
import random
from typing import Callable, List, Optional

import numpy as np
import torch


def set_seed_everywhere(seed: int) -> None:
    """/root/reference/src/utils/pipeline.py:11-22 with train=False."""
    import os

    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    np.random.seed(seed)
    random.seed(seed)
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    os.environ["PYTHONHASHSEED"] = str(seed)


def resolve_adapter_class(env_adapter_path: str):
    """`src.experiments.env_adapters.<simulator>.<Class>` (configuration_pipeline.py:245-251) -> the GPU adapter of the
    same class name in int_act_b200.adapters."""
    from .. import adapters

    name = env_adapter_path.rsplit(".", 1)[-1]
    cls = getattr(adapters, name, None)
    if cls is None:
        raise ValueError(f"no B200 adapter named {name!r} (from {env_adapter_path!r}); available: "
                         "BridgeSimplerAdapter, BridgeSimplerBatchAdapter")
    return cls


class LeRobotPolicyWrapper:
    def __init__(self, model=None, env_adapter=None, *, pipeline_cfg=None, model_class=None,
                 action_step: Optional[int] = None, model_loader: Optional[Callable[[str], object]] = None,
                 adapter_factory: Optional[Callable[[], object]] = None, seed: Optional[int] = None, device=None,
                 max_batch: int = 1):
        """Reference form: `LeRobotPolicyWrapper(pipeline_cfg=cfg, model_class=PI0Policy)` (no model until `switch_model`).
        Direct form: `LeRobotPolicyWrapper(model, env_adapter, action_step=...)` with a ready PI0Policy and adapter;
        `model_loader(path) -> model` / `adapter_factory()` then serve `switch_model`."""
        self.pipeline_cfg = pipeline_cfg
        self.model_class = model_class
        self.model = model
        self.env_adapter = env_adapter
        self._model_loader = model_loader
        self._adapter_factory = adapter_factory
        self.max_batch = max_batch
        # the engine's static buffers and the adapter's device state belong to one caller at a time; host-side work (gathering a
        # group's frames) happens outside this lock, so a second server worker can prepare the next group meanwhile
        self._gpu_lock = threading.RLock()
        if pipeline_cfg is not None:
            if model_class is None:
                raise ValueError("pipeline_cfg given without model_class")
            self.eval_cfg = pipeline_cfg.eval_cfg
            self.model_cfg = pipeline_cfg.model_cfg
            self.action_step = self.eval_cfg.action_step if action_step is None else action_step
            self.port, self.host = self.eval_cfg.port, self.eval_cfg.host
            self.debug = getattr(pipeline_cfg, "debug", False)
            self.seed = pipeline_cfg.seed
            self.gpu_id = getattr(pipeline_cfg, "gpu_id", 0)
            self.device = torch.device(device) if device is not None else torch.device("cuda", int(self.gpu_id))
            self.use_amp = getattr(pipeline_cfg, "use_amp", True)
            self.dtype = torch.bfloat16 if pipeline_cfg.use_bf16 else torch.float32
            set_seed_everywhere(self.seed)
        else:
            if action_step is None:
                if model is None:
                    raise ValueError("action_step is needed when the wrapper starts without a model")
                action_step = model.config.n_action_steps
            self.action_step = action_step
            self.device = torch.device(device) if device is not None else None
            if seed is not None:  # set_seed_everywhere(seed, train=False), policy_wrapper.py:46-47
                torch.manual_seed(seed)
                np.random.seed(seed)

    # ---- control -------------------------------------------------------------------------------------------------
    def _initialze_model_server(self, model_path: str):
        """policy_wrapper.py:86-108 (name kept, typo included): load, `.to(dtype)`, `.to(device)`, `.eval()`."""
        self.model_path = model_path
        cfg = self.pipeline_cfg
        model = self.model_class.from_pretrained(pretrained_name_or_path=model_path, config=self.model_cfg,
                                                 **({"max_batch": self.max_batch} if self.max_batch > 1 else {}))
        if getattr(cfg, "freeze_lm_head", False) and getattr(self.model_cfg, "type", "pi0") == "pi0":
            model.model.paligemma_with_expert.gemma_expert.lm_head = None
            model.model.paligemma_with_expert.paligemma.language_model.lm_head = None
        model = model.to(self.dtype)
        model = model.to(self.device)
        self.model = model  # use_torch_compile is accepted and ignored: there is no tracing compiler on this path
        self.model.eval()

    def _initialize_env_adapter(self):
        """policy_wrapper.py:110-113: the adapter class named by `eval_cfg.env_adapter_path`, built from the pipeline config."""
        return resolve_adapter_class(self.eval_cfg.env_adapter_path)(config=self.pipeline_cfg, device=self.device)

    def switch_model(self, new_model_path):
        """Switch to a new checkpoint: free the old engine, load, rebuild the adapter (policy_wrapper.py:61-67)."""
        self.model = None
        gc.collect()
        if torch.cuda.is_available():
            torch.cuda.empty_cache()
        if self.pipeline_cfg is not None:
            self._initialze_model_server(new_model_path)
            self.env_adapter = self._initialize_env_adapter()
            return
        if self._model_loader is None:
            raise ValueError("switch_model needs a model_loader")
        self.model = self._model_loader(new_model_path)
        if self._adapter_factory is not None:
            self.env_adapter = self._adapter_factory()

    def reset(self):
        """Reset the model's action queue and the adapter (policy_wrapper.py:69-84)."""
        if self.model is None:
            raise ValueError("Model is not initialized. Please initialize model first.")
        if self.env_adapter is None:
            raise ValueError("Env adapter is not initialized. Please initialize env adapter first. env adapter is usually "
                             "initialized when you initialize the model.")
        if hasattr(self.model, "reset") and callable(self.model.reset):
            self.model.reset()
        if hasattr(self.env_adapter, "reset") and callable(self.env_adapter.reset):
            self.env_adapter.reset()

    def warmup(self, max_batch: Optional[int] = None):
        """Builds the engine and captures the CUDA graph of every batch size the server can coalesce (1..max_batch), so
        that no client pays workspace allocation + eager warm-up + graph capture inside its first request."""
        if self.model is not None and hasattr(self.model, "precapture"):
            self.model.precapture(range(1, (max_batch or self.max_batch) + 1))

    # ---- inference -------------------------------------------------------------------------------------------------
    def _require_model(self):
        if self.model is None or self.env_adapter is None:
            raise ValueError("Model is not initialized. Please initialize model first.")

    def select_action(self, element: dict):
        self._require_model()
        with self._gpu_lock:
            element = self.env_adapter.preprocess(element)
            with torch.inference_mode():
                actions = torch.stack([self.model.select_action(element) for _ in range(self.action_step)], dim=1)
            actions = actions[0] if actions.shape[0] == 1 else actions  # no batch dim for a single observation
            return self.env_adapter.postprocess(actions.float())

    def can_batch(self) -> bool:
        """Requests may share an engine batch when every request consumes one whole chunk (no stale queue entries,
        SURVEY.md Q5) and the model exposes the queue-free `infer_action`."""
        return self.model is not None and hasattr(self.model, "infer_action") and \
            self.action_step == self.model.config.n_action_steps

    def select_action_many(self, elements: List[dict]) -> List[np.ndarray]:
        """Several independent requests in one engine batch. Every request is pre-processed by the adapter exactly as
        it would be alone; the noise of the batch is one `torch.normal` draw of shape [n, H, 32] (the reference would
        draw n times [1, H, 32] from the same generator: same distribution, different stream positions).
        A request whose pre-processing fails gets its own exception (an Exception instance in its slot of the returned
        list); the others still run."""
        self._require_model()
        if len(elements) == 1 or not self.can_batch():
            return [self.select_action(e) for e in elements]
        pre, failed, batch, host = [], {}, None, None
        if hasattr(self.env_adapter, "gather_many"):
            try:  # one host concatenation for the whole group, outside the GPU lock; any problem -> per-request path
                sizes = [int(np.asarray(e["observation.state"]).reshape(-1, 8).shape[0]) for e in elements]
                host = self.env_adapter.gather_many(elements)
            except Exception:  # noqa: BLE001
                host = None
        with self._gpu_lock:
            return self._run_many_locked(elements, host, sizes if host is not None else None)

    def _run_many_locked(self, elements, host, sizes):
        pre, failed, batch = [], {}, None
        if host is not None:
            try:  # one H2D + one resize / proprio launch for the whole group
                batch = self.env_adapter.preprocess(host)
                pre = [(i, None) for i in range(len(elements))]
            except Exception:  # noqa: BLE001
                batch, pre = None, []
        elif hasattr(self.env_adapter, "preprocess_many"):
            try:
                sizes = [int(np.asarray(e["observation.state"]).reshape(-1, 8).shape[0]) for e in elements]
                batch = self.env_adapter.preprocess_many(elements)
                pre = [(i, None) for i in range(len(elements))]
            except Exception:  # noqa: BLE001
                batch, pre = None, []
        if batch is None:
            for i, e in enumerate(elements):
                try:
                    pre.append((i, self.env_adapter.preprocess(e)))
                except Exception as exc:  # noqa: BLE001 - reported to the one client that sent it
                    failed[i] = exc
        out: List[object] = [None] * len(elements)
        for i, exc in failed.items():
            out[i] = exc
        if not pre:
            return out
        if batch is None:
            sizes = [p["observation.state"].shape[0] for _, p in pre]
            batch = {}
            for k in pre[0][1]:
                vals = [p[k] for _, p in pre]
                if isinstance(vals[0], torch.Tensor):
                    batch[k] = torch.cat(vals, dim=0)
                else:
                    batch[k] = [t for v in vals for t in (v if isinstance(v, (list, tuple)) else [v])]
        with torch.inference_mode():
            if hasattr(self.model, "reset"):
                self.model.reset()
            actions = self.model.infer_action(batch)[:, : self.action_step]
        # one post-processing launch + one D2H for the whole group (element-wise per action row, so slicing the result gives
        # exactly what per-request calls would); a single-observation request gets its [action_step, 7] array without batch dim
        post = self.env_adapter.postprocess(actions.float() if actions.shape[0] > 1 else actions[0].float())
        if actions.shape[0] == 1:
            post = post[None]
        o = 0
        for (i, _), n in zip(pre, sizes):
            a = post[o:o + n]
            o += n
            out[i] = np.ascontiguousarray(a[0] if n == 1 else a)
        return out
