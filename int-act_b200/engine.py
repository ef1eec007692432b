"""`Pi0Engine`: owns the packed weights, the device workspace and the `pi0_ctx` of one GPU, and runs
`sample_actions` ([mirror] /root/reference/src/model/mvla/modeling_mvla.py:753-804) through the C ABI.

PyTorch is plumbing only (device memory, streams, CUDA-graph capture); all arithmetic happens in libpi0b200.so.
There is no fallback: constructing an engine without the library or without a CUDA device raises.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence

import torch

from . import binding as pb
from .weights import PackedWeights, euler_times


class Pi0Engine:
    def __init__(self, packed: PackedWeights, *, n_cams: int = 1, lang_len: int = 72, horizon: int = 4,
                 num_steps: int = 10, max_batch: int = 1, dtype: str = "bf16", gemm_impl: int = 0,
                 skip_dead_work: bool = True, images_pre_normalized: bool = False, use_cuda_graph: bool = True):
        if not torch.cuda.is_available():
            raise pb.Pi0Error("Pi0Engine needs a CUDA device: the sm_100a library is the only implementation "
                              "of this path (no CPU fallback)")
        self.lib = pb.lib()
        self.packed = packed
        self.device = packed.device
        self.dtype = dtype
        self.act = torch.bfloat16 if dtype == "bf16" else torch.float32
        if packed.act != self.act:
            raise pb.Pi0Error(f"weights packed as {packed.act}, engine dtype {self.act}")
        if packed.num_steps != num_steps:
            raise pb.Pi0Error("weights were packed for a different num_steps (time bias table)")
        # `tokenizer_max_length` is a free config key in the reference (pi0_finetune_bridge.json:46); the kernels want the
        # text block in multiples of 8 tokens. Extra positions are padding with mask 0: masked keys get probability
        # exp(-2.38e38 - max) == 0 exactly and position ids are cumsum(pad) - 1, so the result is bit-identical.
        self.lang_len_user = lang_len
        lang_len = (lang_len + 7) // 8 * 8
        self.n_cams, self.lang_len, self.horizon, self.num_steps = n_cams, lang_len, horizon, num_steps
        self.max_batch = max_batch
        self.P = n_cams * 256 + lang_len
        self.S = 1 + horizon
        self.images_pre_normalized = images_pre_normalized
        props = torch.cuda.get_device_properties(self.device)
        cfg = pb.ModelCfg()
        cfg.n_cams, cfg.lang_len, cfg.horizon = n_cams, lang_len, horizon
        cfg.max_state_dim = cfg.max_action_dim = 32
        cfg.num_steps = num_steps
        cfg.n_euler_iters = len(euler_times(num_steps))
        cfg.dtype = pb.DTYPE_BF16 if dtype == "bf16" else pb.DTYPE_F32
        cfg.gemm_impl = gemm_impl
        cfg.max_batch = max_batch
        cfg.num_sms = props.multi_processor_count
        cfg.skip_dead_work = int(skip_dead_work)
        cfg.n_siglip_layers = packed.dims.n_siglip_layers
        cfg.n_gemma_layers = packed.dims.n_gemma_layers
        cfg.images_pre_normalized = int(images_pre_normalized)
        self.cfg = cfg
        nbytes = self.lib.pi0_workspace_bytes(C.byref(cfg))
        if nbytes == 0:
            pb.check(-2, "pi0_workspace_bytes")
        self.workspace = torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
        self.ctx = C.c_void_p()
        # every call into the library runs with the engine's device current: kernel attributes, cudaMemcpy / cudaMemset of
        # pi0_ctx_create and the launches themselves all act on the CURRENT device of the calling thread
        with torch.cuda.device(self.device):
            pb.check(self.lib.pi0_ctx_create(C.byref(cfg), C.byref(packed.struct), pb.ptr(self.workspace), nbytes,
                                             C.byref(self.ctx)), "pi0_ctx_create")
        self.use_cuda_graph = use_cuda_graph
        self._graphs: Dict[int, dict] = {}
        self.launches_per_infer: Dict[int, int] = {}

    def __del__(self):
        try:
            if getattr(self, "ctx", None):
                self.lib.pi0_ctx_destroy(self.ctx)
                self.ctx = None
        except Exception:
            pass

    # ---- raw call (no graph): all tensors already on the device, contiguous -------------------------------
    def _infer_raw(self, B: int, images: Sequence[torch.Tensor], img_masks: torch.Tensor, tokens: torch.Tensor,
                   lang_masks: torch.Tensor, state: torch.Tensor, x_t: torch.Tensor):
        arr = (C.c_void_p * self.n_cams)(*[t.data_ptr() for t in images])
        scale, shift = (1.0, 0.0) if self.images_pre_normalized else (2.0, -1.0)
        with torch.cuda.device(self.device):
            pb.check(self.lib.pi0_infer(self.ctx, B, arr, scale, shift, pb.ptr(img_masks), pb.ptr(tokens),
                                        pb.ptr(lang_masks), pb.ptr(state), pb.ptr(x_t), pb.current_stream(self.device)),
                     "pi0_infer")

    def _static(self, B: int) -> dict:
        """Static input/output buffers (and, lazily, the captured CUDA graph) for batch size B."""
        st = self._graphs.get(B)
        if st is None:
            dev = self.device
            st = {
                "images": [torch.zeros((B, 3, 224, 224), dtype=torch.float32, device=dev) for _ in range(self.n_cams)],
                "img_masks": torch.ones((self.n_cams, B), dtype=torch.uint8, device=dev),
                "tokens": torch.zeros((B, self.lang_len), dtype=torch.int64, device=dev),
                "lang_masks": torch.zeros((B, self.lang_len), dtype=torch.uint8, device=dev),
                "state": torch.zeros((B, 32), dtype=torch.float32, device=dev),
                "x_t": torch.zeros((B, self.horizon, 32), dtype=torch.float32, device=dev),
                "graph": None,
            }
            self._graphs[B] = st
        return st

    def _run_static(self, B: int):
        with torch.cuda.device(self.device):
            self._run_static_on_device(B)

    def _run_static_on_device(self, B: int):
        st = self._static(B)
        if not self.use_cuda_graph:
            self._infer_raw(B, st["images"], st["img_masks"], st["tokens"], st["lang_masks"], st["state"], st["x_t"])
            return
        if st["graph"] is None:
            # warm-up outside capture (sets kernel attributes, validates arguments), then capture the whole
            # sample_actions launch sequence into one graph: no host work between the ~2k launches at replay
            keep = st["x_t"].clone()
            side = torch.cuda.Stream(device=self.device)
            side.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(side):
                self._infer_raw(B, st["images"], st["img_masks"], st["tokens"], st["lang_masks"], st["state"], st["x_t"])
            torch.cuda.current_stream(self.device).wait_stream(side)
            torch.cuda.synchronize(self.device)
            st["x_t"].copy_(keep)
            g = torch.cuda.CUDAGraph()
            self.lib.pi0_launch_count(1)
            # torch.cuda.graph's default capture stream is created once per process on whatever device was current then:
            # an engine on another GPU must bring its own, or the capture records nothing ("The CUDA Graph is empty")
            with torch.cuda.graph(g, stream=side):
                self._infer_raw(B, st["images"], st["img_masks"], st["tokens"], st["lang_masks"], st["state"], st["x_t"])
            self.launches_per_infer[B] = int(self.lib.pi0_launch_count(1))
            st["x_t"].copy_(keep)
            st["graph"] = g
        st["graph"].replay()

    def warm(self, B: int):
        """Allocates the static buffers of batch size B and captures its CUDA graph on a dummy observation (one valid text
        token, zero image), so the first real request of that size only replays."""
        st = self._static(B)
        if st["graph"] is None:
            st["tokens"][:, 0] = 2
            st["lang_masks"][:, 0] = 1
            self._run_static(B)
            torch.cuda.synchronize(self.device)

    # ---- public: sample_actions -------------------------------------------------------------------------------
    @torch.no_grad()
    def sample_actions(self, images: Sequence[torch.Tensor], img_masks: Sequence[torch.Tensor], lang_tokens: torch.Tensor,
                       lang_masks: torch.Tensor, state: torch.Tensor, noise: Optional[torch.Tensor] = None) -> torch.Tensor:
        """images: n_cams x f32 [B,3,224,224] (range per `images_pre_normalized`), img_masks: n_cams x bool [B],
        lang_tokens int64 [B,L], lang_masks bool [B,L], state [B, <=32], noise f32 [B,H,32] or None.
        Returns x_t f32 [B, H, 32] (a fresh tensor)."""
        B = state.shape[0]
        if B > self.max_batch:
            raise pb.Pi0Error(f"batch {B} > max_batch {self.max_batch}")
        if len(images) != self.n_cams:
            raise pb.Pi0Error(f"{len(images)} cameras given, engine built for {self.n_cams}")
        st = self._static(B)
        dev = self.device
        for dst, src in zip(st["images"], images):
            dst.copy_(src.to(torch.float32), non_blocking=True)
        st["img_masks"].copy_(torch.stack([m.reshape(B) for m in img_masks]).to(torch.uint8), non_blocking=True)
        L = lang_tokens.shape[1]
        if L != self.lang_len_user:
            raise pb.Pi0Error(f"{L} text positions given, engine built for tokenizer_max_length {self.lang_len_user}")
        if L != self.lang_len:
            st["tokens"].zero_()
            st["lang_masks"].zero_()
        st["tokens"][:, :L].copy_(lang_tokens, non_blocking=True)
        st["lang_masks"][:, :L].copy_(lang_masks.to(torch.uint8), non_blocking=True)
        st["state"].zero_()
        st["state"][:, : state.shape[1]].copy_(state.to(torch.float32), non_blocking=True)  # pad_vector
        if noise is None:
            # same generator and call as the reference's sample_noise ([mirror] modeling_mvla.py:567-575)
            noise = torch.normal(mean=0.0, std=1.0, size=(B, self.horizon, 32), dtype=torch.float32, device=dev)
        st["x_t"].copy_(noise.to(torch.float32), non_blocking=True)
        self._run_static(B)
        return st["x_t"].clone()

    def profile_step(self, B: int) -> Dict[str, tuple]:
        """One eager (non-graph) pass over the static buffers of batch B with a CUDA-event pair around every launch.
        Returns {label: (total_ms, launches)}; used by bench.py for the per-kernel roofline."""
        import json

        st = self._static(B)
        keep = st["x_t"].clone()
        torch.cuda.synchronize(self.device)
        pb.check(self.lib.pi0_profile(self.ctx, 1), "pi0_profile")
        try:
            self._infer_raw(B, st["images"], st["img_masks"], st["tokens"], st["lang_masks"], st["state"], st["x_t"])
            buf = C.create_string_buffer(1 << 16)
            pb.check(self.lib.pi0_profile_report(self.ctx, buf, len(buf)), "pi0_profile_report")
        finally:
            self.lib.pi0_profile(self.ctx, 0)
            st["x_t"].copy_(keep)
        return {k: (v[0], v[1]) for k, v in json.loads(buf.value.decode()).items()}

    def path_report(self, B: int) -> Dict[str, object]:
        """Which kernel variant every phase takes at batch size B (fused vs per-op fallbacks made visible)."""
        import json

        buf = C.create_string_buffer(1024)
        pb.check(self.lib.pi0_path_report(self.ctx, B, buf, len(buf)), "pi0_path_report")
        return json.loads(buf.value.decode())

    def check_guards(self) -> None:
        """With PI0_GUARD=1 set before the engine was built: raises if any kernel wrote past the end of a workspace buffer."""
        buf = C.create_string_buffer(512)
        with torch.cuda.device(self.device):
            rc = self.lib.pi0_debug_check_guards(self.ctx, buf, len(buf))
        if rc < 0:
            pb.check(rc, "pi0_debug_check_guards")
        if rc != 0:
            raise pb.Pi0Error(f"workspace guard violated: {buf.value.decode()}")

    def debug_buffer(self, name: str, dtype=None) -> torch.Tensor:
        """Copy of a named workspace buffer (parity tests)."""
        n = C.c_int64()
        p = self.lib.pi0_debug_buffer(self.ctx, name.encode(), C.byref(n))
        if not p:
            raise KeyError(name)
        if dtype is None:
            dtype = {"pad": torch.uint8, "prefix_pos": torch.int32, "suffix_pos": torch.int32,
                     "suffix_cum": torch.int32}.get(name, self.act)
        off = p - self.workspace.data_ptr()
        es = torch.empty((), dtype=dtype).element_size()
        return self.workspace[off: off + n.value * es].view(dtype).clone()
