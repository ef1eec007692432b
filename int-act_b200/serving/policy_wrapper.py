"""`LeRobotPolicyWrapper` for the B200 engine (row a3): the object the policy server calls.

Contract of the reference wrapper (/root/reference/src/experiments/policies/policy_wrapper.py:18-170):
`select_action(obs) -> np.float64[action_step, 7]` (or `[B, action_step, 7]` for a rollout batch),
`reset()`, `switch_model(new_model_path)`. `select_action` = env_adapter.preprocess -> `action_step` calls of
`model.select_action` (queue pops, policy_wrapper.py:131-135) stacked on dim 1 -> drop the batch dim when B == 1 ->
env_adapter.postprocess.

Added for the B200 server: `select_action_many(list of obs)` runs several single-observation requests as ONE engine
batch (pre / post-processing stay per request, so every client gets exactly the adapter it would have got alone).
"""
from __future__ import annotations

from typing import Callable, List, Optional

import numpy as np
import torch


class LeRobotPolicyWrapper:
    def __init__(self, model, env_adapter, *, action_step: Optional[int] = None,
                 model_loader: Optional[Callable[[str], object]] = None,
                 adapter_factory: Optional[Callable[[], object]] = None, seed: Optional[int] = None):
        """model: a PI0Policy (already `.to(dtype).to(device).eval()`); env_adapter: an adapter from
        int_act_b200.adapters; action_step: actions returned per request (default: the model's n_action_steps);
        model_loader(path) -> model: used by `switch_model` (policy_wrapper.py:61-67,86-105)."""
        self.model = model
        self.env_adapter = env_adapter
        self.action_step = action_step if action_step is not None else model.config.n_action_steps
        self._model_loader = model_loader
        self._adapter_factory = adapter_factory
        if seed is not None:  # set_seed_everywhere(seed, train=False), policy_wrapper.py:46-47
            torch.manual_seed(seed)
            np.random.seed(seed)

    # ---- control -------------------------------------------------------------------------------------------------
    def switch_model(self, new_model_path):
        """Switch to a new checkpoint: free the old engine, load, rebuild the adapter (policy_wrapper.py:61-67)."""
        if self._model_loader is None:
            raise ValueError("switch_model needs a model_loader")
        self.model = None
        if torch.cuda.is_available():
            torch.cuda.empty_cache()
        self.model = self._model_loader(new_model_path)
        if self._adapter_factory is not None:
            self.env_adapter = self._adapter_factory()

    def reset(self):
        """Reset the model's action queue and the adapter (policy_wrapper.py:69-84)."""
        if self.model is None:
            raise ValueError("Model is not initialized. Please initialize model first.")
        if self.env_adapter is None:
            raise ValueError("Env adapter is not initialized. Please initialize env adapter first.")
        if hasattr(self.model, "reset") and callable(self.model.reset):
            self.model.reset()
        if hasattr(self.env_adapter, "reset") and callable(self.env_adapter.reset):
            self.env_adapter.reset()

    # ---- inference -------------------------------------------------------------------------------------------------
    def select_action(self, element: dict):
        element = self.env_adapter.preprocess(element)
        with torch.inference_mode():
            actions = torch.stack([self.model.select_action(element) for _ in range(self.action_step)], dim=1)
        actions = actions[0] if actions.shape[0] == 1 else actions  # no batch dim for a single observation
        return self.env_adapter.postprocess(actions.float())

    def can_batch(self) -> bool:
        """Requests may share an engine batch when every request consumes one whole chunk (no stale queue entries,
        SURVEY.md Q5) and the model exposes the queue-free `infer_action`."""
        return hasattr(self.model, "infer_action") and self.action_step == self.model.config.n_action_steps

    def select_action_many(self, elements: List[dict]) -> List[np.ndarray]:
        """Several independent requests in one engine batch. Every request is pre-processed by the adapter exactly as
        it would be alone; the noise of the batch is one `torch.normal` draw of shape [n, H, 32] (the reference would
        draw n times [1, H, 32] from the same generator: same distribution, different stream positions)."""
        if len(elements) == 1 or not self.can_batch():
            return [self.select_action(e) for e in elements]
        pre = [self.env_adapter.preprocess(e) for e in elements]
        sizes = [p["observation.state"].shape[0] for p in pre]
        batch = {}
        for k in pre[0]:
            vals = [p[k] for p in pre]
            if isinstance(vals[0], torch.Tensor):
                batch[k] = torch.cat(vals, dim=0)
            else:
                batch[k] = [t for v in vals for t in (v if isinstance(v, (list, tuple)) else [v])]
        with torch.inference_mode():
            if hasattr(self.model, "reset"):
                self.model.reset()
            actions = self.model.infer_action(batch)[:, : self.action_step]
        out, o = [], 0
        for n in sizes:
            a = actions[o:o + n]
            o += n
            out.append(self.env_adapter.postprocess((a[0] if n == 1 else a).float()))
        return out
