"""`PI0Policy`: the lerobot `PreTrainedPolicy` duck type the reference's server calls, on the B200 engine.

Surface mirrored:
  * `PI0Policy(config, dataset_stats=None)`, `PI0Policy.from_pretrained(pretrained_name_or_path, config=, strict=)`
    (/root/reference/src/experiments/policies/policy_wrapper.py:159-162), `.to(dtype|device)`, `.eval()`, `.reset()`;
  * `.select_action(batch, noise=None) -> Tensor[B, action_dim]` with the `n_action_steps` action queue
    ([mirror] /root/reference/src/model/mvla/modeling_mvla.py:257-259,269-307);
  * `.infer_action(obs, noise=None) -> Tensor[B, n_action_steps, action_dim]`: the chunk `select_action` enqueues
    (the name BASELINE.json's north_star uses);
  * `.model.paligemma_with_expert.{paligemma.language_model,gemma_expert}.lm_head` assignable to None
    (policy_wrapper.py:164-168).
Batch keys: "observation.images.<cam>" f32 [B,3,H,W], "observation.state" [B, state_dim], "task" list[str]
(/root/reference/src/experiments/env_adapters/simpler.py:90-94); optionally pre-tokenised "lang_tokens"/"lang_masks".
"""
from __future__ import annotations

import os
from collections import deque
from types import SimpleNamespace
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

from . import binding as pb
from .config import PI0Config
from .engine import Pi0Engine
from .weights import (Dims, PackedWeights, canonical_state_dict, load_safetensors, random_state_dict, split_checkpoint,
                      stats_from_buffers)

OBS_ROBOT = "observation.state"
ACTION = "action"


def resize_with_pad(img: torch.Tensor, width: int, height: int, pad_value: float = -1) -> torch.Tensor:
    """[mirror] modeling_mvla.py:131-150: bilinear resize keeping the aspect ratio, pad on the left / top."""
    if img.ndim != 4:
        raise ValueError(f"(b,c,h,w) expected, but {img.shape}")
    cur_height, cur_width = img.shape[2:]
    if (cur_height, cur_width) == (height, width):
        return img
    ratio = max(cur_width / width, cur_height / height)
    resized_height = int(cur_height / ratio)
    resized_width = int(cur_width / ratio)
    resized = F.interpolate(img, size=(resized_height, resized_width), mode="bilinear", align_corners=False)
    pad_height = max(0, int(height - resized_height))
    pad_width = max(0, int(width - resized_width))
    return F.pad(resized, (pad_width, 0, pad_height, 0), value=pad_value)


class SyntheticTokenizer:
    """Deterministic stand-in used ONLY where the PaliGemma SentencePiece files are unavailable (no network on this
    box): bos=2, one id per whitespace word hashed into [3, vocab), "\\n" -> 108, right-padded with 0.
    It preserves shapes/masks of the real tokenizer call ([mirror] modeling_mvla.py:387-407), not its ids."""

    def __init__(self, vocab: int = 257152):
        self.vocab = vocab

    def __call__(self, tasks: List[str], max_length: int):
        ids = torch.zeros((len(tasks), max_length), dtype=torch.int64)
        mask = torch.zeros((len(tasks), max_length), dtype=torch.bool)
        for i, t in enumerate(tasks):
            toks = [2]
            for wd in t.rstrip("\n").split():
                h = 2166136261
                for ch in wd.encode():
                    h = ((h ^ ch) * 16777619) & 0xFFFFFFFF
                toks.append(3 + h % (self.vocab - 3))
            toks.append(108)
            toks = toks[:max_length]  # truncation=True
            ids[i, : len(toks)] = torch.tensor(toks)
            mask[i, : len(toks)] = True
        return ids, mask


class HFTokenizer:
    def __init__(self, name="google/paligemma-3b-pt-224"):
        from transformers import AutoTokenizer

        self.tok = AutoTokenizer.from_pretrained(name, local_files_only=True)

    def __call__(self, tasks, max_length):
        out = self.tok(tasks, padding="max_length", padding_side="right", max_length=max_length, return_tensors="pt",
                       truncation=True)
        return out["input_ids"], out["attention_mask"].bool()


# ---- adapt_to_pi_aloha ([mirror] modeling_mvla.py:166-218,409-425): joint flips + gripper linear <-> angular maps ------
def _aloha_gripper_to_angular(value: torch.Tensor) -> torch.Tensor:
    value = value * (0.05800 - 0.01844) + 0.01844               # Aloha's linear gripper position
    horn_radius, arm_length = 0.022, 0.036                      # Interbotix linkage constants
    value = (horn_radius ** 2 + value ** 2 - arm_length ** 2) / (2 * horn_radius * value)
    value = torch.arcsin(torch.clamp(value, -1.0, 1.0))
    return (value - 0.4) / (1.5 - 0.4)                          # measured joint range -> [0, 1]


def _aloha_gripper_from_angular(value: torch.Tensor) -> torch.Tensor:
    value = value * (1.5 - 0.4) + 0.4
    return (value - (-0.6213)) / (1.4910 - (-0.6213))


def pi_aloha_decode_state(state: torch.Tensor) -> torch.Tensor:
    state = state.clone()
    for motor_idx in (1, 2, 8, 9):
        state[:, motor_idx] *= -1
    for motor_idx in (6, 13):
        state[:, motor_idx] = _aloha_gripper_to_angular(state[:, motor_idx])
    return state


def pi_aloha_encode_actions(actions: torch.Tensor) -> torch.Tensor:
    actions = actions.clone()
    for motor_idx in (1, 2, 8, 9):
        actions[:, :, motor_idx] *= -1
    for motor_idx in (6, 13):
        actions[:, :, motor_idx] = _aloha_gripper_from_angular(actions[:, :, motor_idx])
    return actions


class _Normalizer:
    """lerobot Normalize / Unnormalize restated: IDENTITY, MEAN_STD ((x-mean)/(std+1e-8)), MIN_MAX (to [-1,1])."""

    def __init__(self, features, mapping, stats, inverse):
        self.items = []
        for key, ft in features.items():
            mode = mapping.get(ft.type, "IDENTITY")
            if mode == "IDENTITY":
                continue
            if stats is None or key not in stats:
                raise ValueError(f"normalization {mode} for {key!r} needs dataset_stats[{key!r}]")
            self.items.append((key, mode, {k: torch.as_tensor(v, dtype=torch.float32) for k, v in stats[key].items()}))
        self.inverse = inverse

    def __call__(self, batch: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
        if not self.items:
            return batch
        batch = dict(batch)
        for key, mode, st in self.items:
            if key not in batch:
                continue
            x = batch[key]
            st = {k: v.to(x.device) for k, v in st.items()}
            if mode == "MEAN_STD":
                batch[key] = x * st["std"] + st["mean"] if self.inverse else (x - st["mean"]) / (st["std"] + 1e-8)
            elif mode == "MIN_MAX":
                if self.inverse:
                    batch[key] = (x + 1) / 2 * (st["max"] - st["min"]) + st["min"]
                else:
                    batch[key] = (x - st["min"]) / (st["max"] - st["min"] + 1e-8) * 2 - 1
            else:
                raise ValueError(mode)
        return batch


class PI0Policy:
    name = "pi0"
    config_class = PI0Config

    def __init__(self, config: PI0Config, dataset_stats: Optional[dict] = None, *, state_dict: Optional[dict] = None,
                 dims: Optional[Dims] = None, seed: int = 0, max_batch: int = 1, tokenizer=None,
                 use_cuda_graph: bool = True, device=None):
        config.validate_features()
        self.config = config
        self.normalize_inputs = _Normalizer(config.input_features, config.normalization_mapping, dataset_stats, False)
        self.unnormalize_outputs = _Normalizer(config.output_features, config.normalization_mapping, dataset_stats, True)
        self._tokenizer = tokenizer
        self._state_dict = state_dict
        self._dims = dims or Dims()
        self._seed = seed
        self._dtype = torch.bfloat16
        self._device = torch.device(device) if device is not None else None
        self._max_batch = max_batch
        self._use_cuda_graph = use_cuda_graph
        self._engine: Optional[Pi0Engine] = None
        self.training = False
        # attribute shims for callers that null the LM heads (policy_wrapper.py:164-168; trainer.py:110,115)
        self.model = SimpleNamespace(paligemma_with_expert=SimpleNamespace(
            paligemma=SimpleNamespace(language_model=SimpleNamespace(lm_head=None)),
            gemma_expert=SimpleNamespace(lm_head=None)))
        self.reset()

    # ---- lerobot PreTrainedPolicy surface ------------------------------------------------------------------
    @classmethod
    def from_pretrained(cls, pretrained_name_or_path: str, *, config: Optional[PI0Config] = None, strict: bool = False,
                        **kw) -> "PI0Policy":
        if config is None:
            config = PI0Config.from_pretrained(pretrained_name_or_path)
        config.validate_features()
        raw = load_safetensors(pretrained_name_or_path, device="cpu")
        sd = canonical_state_dict(raw, strict=strict)  # key / shape validation, tied lm_head alias
        stats = kw.pop("dataset_stats", None)
        if stats is None:
            # lerobot keeps the dataset statistics inside the checkpoint as Normalize / Unnormalize buffers: the reference
            # wrapper never passes dataset_stats (policy_wrapper.py:159-162) and MEAN_STD configs (hf_pi0.json:4-8) rely on them
            stats = stats_from_buffers(split_checkpoint(raw)[1], config.input_features, config.output_features) or None
        return cls(config, stats, state_dict=sd, **kw)

    def to(self, arg):
        if isinstance(arg, torch.dtype):
            if arg not in (torch.bfloat16, torch.float32):
                raise ValueError("PI0Policy supports bfloat16 (tcgen05 product path) and float32 (check mode)")
            if arg != self._dtype:
                self._dtype = arg
                self._engine = None
        else:
            dev = torch.device(arg)
            if dev.type != "cuda":
                if dev.type == "cpu" and self._engine is None:
                    return self  # configs carry device="cpu" until the wrapper moves the model (policy_wrapper.py:94)
                raise pb.Pi0Error("PI0Policy runs on CUDA only (no CPU fallback)")
            if dev != self._device:
                self._device = dev
                self._engine = None
        return self

    def eval(self):
        self.training = False
        return self

    def train(self, mode: bool = True):
        if mode:
            raise NotImplementedError("the B200 path is inference-only")
        return self.eval()

    def parameters(self):
        eng = self._ensure_engine()
        return iter(eng.packed.keep)

    def reset(self):
        """This should be called whenever the environment is reset ([mirror] modeling_mvla.py:257-259)."""
        self._action_queue = deque([], maxlen=self.config.n_action_steps)

    # ---- engine ------------------------------------------------------------------------------------------------
    def _ensure_engine(self, batch_size: int = 1) -> Pi0Engine:
        if self._engine is not None and batch_size <= self._engine.max_batch:
            return self._engine
        if self._device is None:
            self._device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else None
        if self._device is None:
            raise pb.Pi0Error("PI0Policy needs a CUDA device (no CPU fallback)")
        packed = self._engine.packed if self._engine is not None else None
        if packed is None:
            sd = self._state_dict
            if sd is None:
                # paligemma_pretrained_path None => random init from config
                sd = random_state_dict(self._dims, device=self._device, seed=self._seed)
            packed = PackedWeights(sd, self._device, self._dtype, num_steps=self.config.num_steps)
        self._max_batch = max(self._max_batch, batch_size)
        self._engine = Pi0Engine(packed, n_cams=self.config.n_cams, lang_len=self.config.tokenizer_max_length,
                                 horizon=self.config.n_action_steps, num_steps=self.config.num_steps,
                                 max_batch=self._max_batch, dtype="bf16" if self._dtype == torch.bfloat16 else "fp32",
                                 images_pre_normalized=self.config.images_pre_normalized,
                                 use_cuda_graph=self._use_cuda_graph)
        return self._engine

    def precapture(self, batch_sizes):
        """Builds the engine for the largest batch size and captures the CUDA graph of each size now (server start-up),
        instead of inside the latency of the first request that happens to have that size."""
        sizes = sorted({int(b) for b in batch_sizes if int(b) >= 1})
        if not sizes:
            return
        eng = self._ensure_engine(sizes[-1])
        for b in sizes:
            eng.warm(b)

    # ---- preparation ([mirror] modeling_mvla.py:344-407,436-439) ------------------------------------------------------
    def prepare_images(self, batch):
        images, img_masks = [], []
        present = [k for k in self.config.image_features if k in batch]
        missing = [k for k in self.config.image_features if k not in batch]
        if len(present) == 0:
            raise ValueError(f"All image features are missing from the batch. At least one expected. "
                             f"(batch: {batch.keys()}) (image_features:{self.config.image_features})")
        img = mask = None
        for key in present:
            img = batch[key]
            if self.config.resize_imgs_with_padding is not None:
                img = resize_with_pad(img, *self.config.resize_imgs_with_padding, pad_value=0)
            # "img * 2 - 1" is fused into the patch im2col kernel (scale/shift of pi0_infer)
            mask = torch.ones(img.shape[0], dtype=torch.bool, device=img.device)
            images.append(img)
            img_masks.append(mask)
        fill = -1.0 if self.config.images_pre_normalized else 0.0  # the kernel maps 0 -> -1 via img*2-1
        for n, _ in enumerate(missing):
            if n >= self.config.empty_cameras:
                break
            images.append(torch.full_like(img, fill))
            img_masks.append(torch.zeros_like(mask))
        return images, img_masks

    def prepare_language(self, batch):
        if "lang_tokens" in batch:
            return batch["lang_tokens"], batch["lang_masks"].bool()
        tasks = batch["task"]
        if isinstance(tasks, str):
            tasks = [tasks]
        tasks = [t if t.endswith("\n") else f"{t}\n" for t in tasks]
        if self._tokenizer is None:
            try:
                self._tokenizer = HFTokenizer()
            except Exception as e:  # tokenizer files are not on this box
                raise pb.Pi0Error(
                    "PaliGemma tokenizer files unavailable; pass tokenizer=SyntheticTokenizer() for synthetic runs or "
                    "provide batch['lang_tokens'/'lang_masks']") from e
        return self._tokenizer(tasks, self.config.tokenizer_max_length)

    # ---- inference -------------------------------------------------------------------------------------------------
    @torch.no_grad()
    def infer_action(self, batch: Dict[str, torch.Tensor], noise: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One action chunk per observation: Tensor[B, n_action_steps, action_dim] (what select_action enqueues)."""
        if self.config.adapt_to_pi_aloha:
            batch = dict(batch)
            batch[OBS_ROBOT] = pi_aloha_decode_state(batch[OBS_ROBOT])
        batch = self.normalize_inputs(batch)
        images, img_masks = self.prepare_images(batch)
        state = batch[OBS_ROBOT]
        lang_tokens, lang_masks = self.prepare_language(batch)
        eng = self._ensure_engine(state.shape[0])
        actions = eng.sample_actions(images, img_masks, lang_tokens, lang_masks, state, noise)
        original_action_dim = self.config.action_feature.shape[0]
        actions = actions[:, :, :original_action_dim]
        actions = self.unnormalize_outputs({ACTION: actions})[ACTION]
        if self.config.adapt_to_pi_aloha:
            actions = pi_aloha_encode_actions(actions)
        return actions

    @torch.no_grad()
    def select_action(self, batch: Dict[str, torch.Tensor], noise: Optional[torch.Tensor] = None) -> torch.Tensor:
        self.eval()
        if len(self._action_queue) == 0:
            actions = self.infer_action(batch, noise)
            self._action_queue.extend(actions.transpose(0, 1))
        return self._action_queue.popleft()
