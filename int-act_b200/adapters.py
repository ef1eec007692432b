"""Environment adapters of the Bridge / SimplerEnv evaluation, B200 side.

Same class names, constructor argument and `preprocess` / `postprocess` / `reset` contract as the reference's
`src/experiments/env_adapters/simpler.py` (BridgeSimplerAdapter) and `simplerMS3.py` (BridgeSimplerBatchAdapter), so a
`LeRobotPolicyWrapper` can hold either implementation. What changes is where the work happens: the rollout batch is
uploaded ONCE as raw uint8 frames + eef poses, and resize / normalisation / quaternion -> euler / bound normalisation
(pre) and de-normalisation / euler -> axis-angle / gripper binarisation (post) run as CUDA kernels
(csrc/adapters.cu, C ABI `pi0_adapter_*`) instead of a per-sample Python loop on the host.

There is no CPU fallback for the batch adapter: without the library / a CUDA device construction fails.
The single-observation adapter resizes with a CUDA restatement of OpenCV's uint8 INTER_LANCZOS4 (8-tap separable,
11-bit fixed-point coefficients; bit-exact against cv2 itself in the tests); its proprio / post-processing use the same
kernels with B = 1. Neither adapter needs OpenCV at run time.
"""
from __future__ import annotations

import json
import math
import warnings
from types import SimpleNamespace
from typing import Optional, Sequence

import numpy as np
import torch

from . import binding as pb

# EE pose in Bridge data is relative to a top-down pose (simpler.py:160-162); baked into the proprio kernel
DEFAULT_ROT = np.array([[0, 0, 1.0], [0, 1.0, 0], [-1.0, 0, 0]])


def u8_bilinear_tables(in_size: int, out_size: int):
    """Per-axis tables of torch's uint8 bilinear resize (align_corners=False, antialias=False), the op
    simplerMS3.py:55-58 applies: source indices and interpolation weights in double precision, weights converted to
    int16 with the largest precision whose maximum still fits 15 bits. idx int32 [out, 2], w int16 [out, 2], precision."""
    scale = float(in_size) / float(out_size)
    idx = np.zeros((out_size, 2), np.int32)
    w = np.zeros((out_size, 2), np.float64)
    for i in range(out_size):
        real = max(scale * (i + 0.5) - 0.5, 0.0)
        i0 = min(int(math.floor(real)), in_size - 1)
        lam = min(max(real - i0, 0.0), 1.0)
        idx[i, 0], idx[i, 1] = i0, i0 + (1 if i0 < in_size - 1 else 0)
        w[i, 0], w[i, 1] = 1.0 - lam, lam
    prec = 0
    for prec in range(22):
        if int(0.5 + w.max() * (1 << (prec + 1))) >= (1 << 15):
            break
    wi = np.where(w * (1 << prec) < 0, w * (1 << prec) - 0.5, w * (1 << prec) + 0.5).astype(np.int64).astype(np.int16)
    return idx, wi, prec


def lanczos4_tables(in_size: int, out_size: int):
    """Per-axis tables of OpenCV's INTER_LANCZOS4 resize on uint8 (the call of simpler.py:47-51; imgproc resize.cpp):
    source offset = floor(fx) with fx = float32((d + 0.5) * scale - 0.5), 8 float32 coefficients from
    interpolateLanczos4 (sin / cos of -(x + 3 - i) pi / 4 over y^2, normalised), converted to 11-bit fixed point with
    saturate_cast<short> (round half to even). ofs int32 [out], coef int16 [out, 8]."""
    s45 = 0.70710678118654752440084436210485
    cs = ((1, 0), (-s45, -s45), (0, 1), (s45, -s45), (-1, 0), (s45, s45), (0, -1), (-s45, s45))
    scale = float(in_size) / out_size
    ofs = np.zeros(out_size, np.int32)
    coef = np.zeros((out_size, 8), np.int16)
    for d in range(out_size):
        f = np.float32((d + 0.5) * scale - 0.5)
        sx = int(math.floor(f))
        x = np.float32(f - np.float32(sx))
        c = np.zeros(8, np.float32)
        if x < np.finfo(np.float32).eps:
            c[3] = 1.0
        else:
            y0 = -(float(x) + 3) * math.pi * 0.25
            s0, c0 = math.sin(y0), math.cos(y0)
            total = np.float32(0)
            for i in range(8):
                y = -(float(x) + 3 - i) * math.pi * 0.25
                c[i] = np.float32((cs[i][0] * s0 + cs[i][1] * c0) / (y * y))
                total = np.float32(total + c[i])
            c = (c * (np.float32(1.0) / total)).astype(np.float32)
        ofs[d] = sx
        coef[d] = np.clip(np.rint(c * np.float32(2048)), -32768, 32767).astype(np.int16)
    return ofs, coef


def _cfg_get(config, path: Sequence[str], default=None):
    cur = config
    for name in path:
        if cur is None:
            return default
        cur = cur.get(name, None) if isinstance(cur, dict) else getattr(cur, name, None)
    return default if cur is None else cur


def _on_device(fn):
    """Runs an adapter method with the adapter's device current (the calling thread may sit on another GPU)."""
    import functools

    @functools.wraps(fn)
    def wrapped(self, *a, **kw):
        with torch.cuda.device(self.device):
            return fn(self, *a, **kw)
    return wrapped


class _GpuAdapterBase:
    """Shared state: statistics on the device (float64), resize tables per input size, the C-ABI handle."""

    def __init__(self, config=None, *, device=None, dataset_statistics: Optional[dict] = None, **kwargs):
        if config is None:
            config = kwargs.pop("config", None)
        if not torch.cuda.is_available():
            raise pb.Pi0Error("the GPU environment adapters need a CUDA device (no CPU fallback)")
        self.lib = pb.lib()
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.image_size = tuple(_cfg_get(config, ("env", "image_size"), (224, 224)))
        self.action_normalization_type = _cfg_get(config, ("env", "action_normalization_type"), "bound")
        self.state_normalization_type = _cfg_get(config, ("env", "state_normalization_type"), "bound")
        assert self.action_normalization_type in ["bound", "gaussian"]
        assert self.state_normalization_type in ["bound", "gaussian"]
        if dataset_statistics is None:
            with open(_cfg_get(config, ("env", "dataset_statistics_path")), "r") as f:
                dataset_statistics = json.load(f)
        self.dataset_statistics = dataset_statistics
        self.dtype = torch.bfloat16 if _cfg_get(config, ("use_bf16",), False) else torch.float32
        self.seed = _cfg_get(config, ("seed",), 42)

        def dev64(v):
            return torch.tensor(np.asarray(v, dtype=np.float64), dtype=torch.float64, device=self.device)

        keys = ("p01", "p99") if self.state_normalization_type == "bound" else ("mean", "std")
        self._pa, self._pb = dev64(dataset_statistics["proprio"][keys[0]]), dev64(dataset_statistics["proprio"][keys[1]])
        keys = ("p01", "p99") if self.action_normalization_type == "bound" else ("mean", "std")
        self._aa, self._ab = dev64(dataset_statistics["action"][keys[0]]), dev64(dataset_statistics["action"][keys[1]])
        self._tables = {}

    def reset(self):
        pass

    # ---- kernels ---------------------------------------------------------------------------------------------------
    def _axis_tables(self, in_size: int, out_size: int):
        key = (in_size, out_size)
        if key not in self._tables:
            idx, w, prec = u8_bilinear_tables(in_size, out_size)
            self._tables[key] = (torch.from_numpy(idx).to(self.device), torch.from_numpy(w).to(self.device), prec)
        return self._tables[key]

    @_on_device
    def resize_normalize(self, images_u8: torch.Tensor) -> torch.Tensor:
        """uint8 [B, H, W, 3] on the device -> float32 [B, 3, h, w] in [-1, 1] (simplerMS3.py:50-63)."""
        B, H, W, C = images_u8.shape
        if C != 3 or images_u8.dtype != torch.uint8:
            raise ValueError(f"expected uint8 [B, H, W, 3] frames, got {images_u8.dtype} {tuple(images_u8.shape)}")
        oh, ow = self.image_size
        xi, xw, xp = self._axis_tables(W, ow)
        yi, yw, yp = self._axis_tables(H, oh)
        out = torch.empty((B, 3, oh, ow), dtype=torch.float32, device=self.device)
        images_u8 = images_u8.contiguous()
        pb.check(self.lib.pi0_adapter_resize_normalize(pb.ptr(images_u8), pb.ptr(out), B, H, W, oh, ow, pb.ptr(xi),
                                                       pb.ptr(xw), xp, pb.ptr(yi), pb.ptr(yw), yp, 1.0 / 255.0, 0.5, 0.5,
                                                       pb.current_stream(self.device)), "pi0_adapter_resize_normalize")
        return out

    @_on_device
    def lanczos4_normalize(self, images_u8: torch.Tensor) -> torch.Tensor:
        """uint8 [B, H, W, 3] on the device -> float32 [B, 3, h, w] in [-1, 1]: cv2.INTER_LANCZOS4 resize (bit-exact with
        OpenCV's generic uint8 path) + x/255 + (x-0.5)/0.5 (simpler.py:47-65)."""
        B, H, W, C = images_u8.shape
        if C != 3 or images_u8.dtype != torch.uint8:
            raise ValueError(f"expected uint8 [B, H, W, 3] frames, got {images_u8.dtype} {tuple(images_u8.shape)}")
        ow, oh = self.image_size  # cv2.resize takes dsize = (width, height)
        key = ("lanczos4", H, W, oh, ow)
        if key not in self._tables:
            xo, xc = lanczos4_tables(W, ow)
            yo, yc = lanczos4_tables(H, oh)
            self._tables[key] = tuple(torch.from_numpy(t).to(self.device) for t in (xo, xc, yo, yc))
        xo, xc, yo, yc = self._tables[key]
        out = torch.empty((B, 3, oh, ow), dtype=torch.float32, device=self.device)
        images_u8 = images_u8.contiguous()
        pb.check(self.lib.pi0_adapter_lanczos4_normalize(pb.ptr(images_u8), pb.ptr(out), B, H, W, oh, ow, pb.ptr(xo),
                                                         pb.ptr(xc), pb.ptr(yo), pb.ptr(yc), 1.0 / 255.0, 0.5, 0.5,
                                                         pb.current_stream(self.device)), "pi0_adapter_lanczos4_normalize")
        return out

    @_on_device
    def proprio(self, eef: torch.Tensor) -> torch.Tensor:
        """eef [B, 8] (float32 / float64, device) -> normalised state float32 [B, 7] carrying the rounding of
        `torch.as_tensor(proprio, dtype=self.dtype)` (simplerMS3.py:66-88,153-165)."""
        if eef.dtype not in (torch.float32, torch.float64):
            eef = eef.to(torch.float64)
        eef = eef.contiguous()
        B = eef.shape[0]
        out = torch.empty((B, 7), dtype=torch.float32, device=self.device)
        pb.check(self.lib.pi0_adapter_proprio(pb.ptr(eef), int(eef.dtype == torch.float64), pb.ptr(out), B,
                                              pb.ptr(self._pa), pb.ptr(self._pb),
                                              0 if self.state_normalization_type == "bound" else 1,
                                              int(self.dtype == torch.bfloat16), pb.current_stream(self.device)),
                 "pi0_adapter_proprio")
        return out

    @_on_device
    def _postprocess_device(self, actions: torch.Tensor) -> torch.Tensor:
        a = actions.to(self.device, torch.float32).contiguous()
        n = a.numel() // 7
        out = torch.empty(a.shape, dtype=torch.float64, device=self.device)
        pb.check(self.lib.pi0_adapter_postprocess(pb.ptr(a), pb.ptr(out), n, pb.ptr(self._aa), pb.ptr(self._ab),
                                                  0 if self.action_normalization_type == "bound" else 1,
                                                  pb.current_stream(self.device)), "pi0_adapter_postprocess")
        return out

    def postprocess(self, actions) -> np.ndarray:
        """actions [action_step, 7] or [B, action_step, 7] (tensor on any device, or ndarray) -> float64 ndarray of the
        same shape: the array the policy server sends back (simpler.py:96-139; simplerMS3.py:95-132)."""
        if isinstance(actions, np.ndarray):
            actions = torch.from_numpy(np.ascontiguousarray(actions, dtype=np.float32))
        return self._postprocess_device(actions).cpu().numpy()


class BridgeSimplerBatchAdapter(_GpuAdapterBase):
    """Batched adapter (reference: simplerMS3.py:135-172). obs = {"observation.images.top": uint8 [B, H, W, 3],
    "observation.state": eef [B, 8], "task": list[str]}."""

    def preprocess(self, obs: dict) -> dict:
        frames = obs["observation.images.top"]
        with warnings.catch_warnings():  # frames unpacked from the wire are read-only views; they are only read here
            warnings.simplefilter("ignore", UserWarning)
            frames = torch.as_tensor(frames)
        if frames.dtype != torch.uint8:
            frames = frames.to(torch.uint8)  # torch.as_tensor(..., dtype=torch.uint8), simplerMS3.py:50
        images = self.resize_normalize(frames.to(self.device, non_blocking=True))
        eef = torch.as_tensor(np.asarray(obs["observation.state"])) if not torch.is_tensor(obs["observation.state"]) \
            else obs["observation.state"]
        state = self.proprio(eef.to(self.device, non_blocking=True))
        return {"observation.images.top": images, "observation.state": state, "task": obs["task"]}


    @staticmethod
    def gather_many(obs_list) -> dict:
        """HOST half of `preprocess_many`: several requests (each a [b_i, H, W, 3] frame stack) concatenated into one
        observation dict. Touches no GPU state, so the server can run it for the next group while the current one computes."""
        frames = np.concatenate([np.asarray(o["observation.images.top"]) for o in obs_list], axis=0)
        eef = np.concatenate([np.asarray(o["observation.state"], dtype=np.float32).reshape(-1, 8) for o in obs_list], axis=0)
        tasks = [t for o in obs_list for t in ([o["task"]] if isinstance(o["task"], str) else list(o["task"]))]
        if frames.shape[0] != eef.shape[0] or frames.shape[0] != len(tasks):
            raise ValueError("frames / states / tasks disagree on the number of observations")
        return {"observation.images.top": frames, "observation.state": eef, "task": tasks}

    def preprocess_many(self, obs_list) -> dict:
        """Several requests as ONE host concatenation, one H2D copy and one resize / proprio launch - the coalescing server's
        path. Same arithmetic per observation as `preprocess`."""
        return self.preprocess(self.gather_many(obs_list))


class BridgeSimplerAdapter(_GpuAdapterBase):
    """Single-observation adapter (reference: simpler.py:154-190). obs = {"observation.images.top": uint8 [H, W, 3],
    "observation.state": the simulator's nested obs dict (eef pose under ["agent"]["eef_pos"]), "task": str}."""

    def preprocess(self, obs: dict) -> dict:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore", UserWarning)
            frame = torch.as_tensor(np.asarray(obs["observation.images.top"]))
        if frame.dtype != torch.uint8:
            frame = frame.to(torch.uint8)
        # cv2.resize(..., INTER_LANCZOS4) + process_images as one kernel on the raw frame (simpler.py:47-65)
        images = self.lanczos4_normalize(frame[None].to(self.device, non_blocking=True))
        eef = np.asarray(obs["observation.state"]["agent"]["eef_pos"])
        state = self.proprio(torch.as_tensor(eef)[None].to(self.device, non_blocking=True))
        return {"observation.images.top": images, "observation.state": state, "task": [obs["task"]]}


def adapter_config(dataset_statistics_path: str, *, image_size=(224, 224), use_bf16: bool = True, seed: int = 42,
                   action_normalization_type: str = "bound", state_normalization_type: str = "bound"):
    """A minimal stand-in for the reference's pipeline config (the attributes the adapters read)."""
    return SimpleNamespace(env=SimpleNamespace(image_size=list(image_size),
                                               action_normalization_type=action_normalization_type,
                                               state_normalization_type=state_normalization_type,
                                               dataset_statistics_path=dataset_statistics_path),
                           use_bf16=use_bf16, seed=seed)
