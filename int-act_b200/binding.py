"""ctypes binding of libpi0b200.so (C ABI declared in include/pi0b200.h).

The library is the product path: there is no CPU or PyTorch fallback. `lib()` raises if the shared
object is missing (build it with `python int-act_b200/build.py`); compute entry points additionally need a
CUDA device. Tensors are passed as raw device pointers (`tensor.data_ptr()`) plus explicit sizes.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PI0_LIB") or os.path.join(_HERE, "libpi0b200.so")  # PI0_LIB: instrumented dev builds

SIGLIP_LAYERS = 27
GEMMA_LAYERS = 18
ABI_VERSION = 4

c_f32p = C.c_void_p  # device pointers are opaque to ctypes
c_ptr = C.c_void_p


class ModelCfg(C.Structure):
    _fields_ = [
        ("n_cams", C.c_int32), ("lang_len", C.c_int32), ("horizon", C.c_int32),
        ("max_state_dim", C.c_int32), ("max_action_dim", C.c_int32), ("num_steps", C.c_int32),
        ("dtype", C.c_int32), ("gemm_impl", C.c_int32), ("max_batch", C.c_int32), ("num_sms", C.c_int32),
        ("skip_dead_work", C.c_int32), ("n_siglip_layers", C.c_int32), ("n_gemma_layers", C.c_int32),
        ("images_pre_normalized", C.c_int32), ("n_euler_iters", C.c_int32), ("reserved", C.c_int32),
    ]


class SiglipLayerW(C.Structure):
    _fields_ = [(n, c_ptr) for n in (
        "ln1_w", "ln1_b", "wqkv", "bqkv", "wo", "bo", "ln2_w", "ln2_b", "wfc1", "bfc1", "wfc2", "bfc2")]


class GemmaLayerW(C.Structure):
    _fields_ = [(n, c_ptr) for n in ("g1", "wqkv", "wo", "g2", "wgu", "wdown")]


class Weights(C.Structure):
    _fields_ = [
        ("patch_w", c_ptr), ("patch_b", c_ptr), ("pos_emb", c_ptr),
        ("sig", SiglipLayerW * SIGLIP_LAYERS),
        ("post_ln_w", c_ptr), ("post_ln_b", c_ptr), ("proj_w", c_ptr), ("proj_b", c_ptr),
        ("embed", c_ptr), ("vocab", C.c_int64),
        ("gem", GemmaLayerW * GEMMA_LAYERS),
        ("exp", GemmaLayerW * GEMMA_LAYERS),
        ("exp_norm_g", c_ptr),
        ("state_w", c_ptr), ("state_b", c_ptr), ("action_in_w", c_ptr), ("action_in_b", c_ptr),
        ("time_in_wa", c_ptr), ("time_bias", c_ptr), ("time_out_w", c_ptr), ("time_out_b", c_ptr),
        ("action_out_w", c_ptr), ("action_out_b", c_ptr),
        ("rope_cos", c_ptr), ("rope_sin", c_ptr), ("max_pos", C.c_int32), ("reserved", C.c_int32),
    ]


class GemmDesc(C.Structure):
    _fields_ = [
        ("A", c_ptr), ("lda", C.c_int64), ("a_z0", C.c_int64), ("a_z1", C.c_int64),
        ("W", c_ptr), ("ldw", C.c_int64), ("w_z0", C.c_int64), ("w_z1", C.c_int64),
        ("M", C.c_int32), ("N", C.c_int32), ("K", C.c_int32), ("nz0", C.c_int32), ("nz1", C.c_int32),
        ("epi_mode", C.c_int32), ("bn", C.c_int32), ("act", C.c_int32),
        ("out", c_ptr), ("ldc", C.c_int64), ("out_z0", C.c_int64), ("out_z1", C.c_int64),
        ("bias", c_ptr),
        ("res", c_ptr), ("ldr", C.c_int64), ("res_row_mod", C.c_int32),
        ("alpha", C.c_float), ("post_div", C.c_float), ("post_mul", C.c_float),
        ("rpb_in", C.c_int32), ("rpb_out", C.c_int32), ("row_off", C.c_int32),
        ("vt", c_ptr), ("vt_col0", C.c_int32), ("vt_hd", C.c_int32), ("vt_heads", C.c_int32),
        ("vt_rpb", C.c_int32), ("vt_toff", C.c_int32), ("vt_pitch", C.c_int64),
        ("rope_cos", c_ptr), ("rope_sin", c_ptr), ("pos_ids", c_ptr),
        ("kcache", c_ptr), ("n_q_heads", C.c_int32), ("kv_pitch", C.c_int64),
        ("dt", C.c_float), ("splits", C.c_int32),
        ("rowsq_out", c_ptr), ("rowsq_ld", C.c_int32),
        ("norm_gain", c_ptr), ("norm_rowsq", c_ptr), ("norm_slots", C.c_int32), ("norm_eps", C.c_float),
    ]


EPI_STORE, EPI_GEGLU, EPI_ROPE, EPI_EULER = 0, 1, 2, 3
ACT_NONE, ACT_GELU_TANH, ACT_SILU = 0, 1, 2
DTYPE_BF16, DTYPE_F32 = 0, 1

# name -> (restype, argtypes); every symbol include/pi0b200.h declares
SIGNATURES = {
    "pi0_abi_version": (C.c_int, []),
    "pi0_last_error": (C.c_char_p, []),
    "pi0_workspace_bytes": (C.c_size_t, [C.POINTER(ModelCfg)]),
    "pi0_ctx_create": (C.c_int, [C.POINTER(ModelCfg), C.POINTER(Weights), c_ptr, C.c_size_t, C.POINTER(c_ptr)]),
    "pi0_ctx_destroy": (None, [c_ptr]),
    "pi0_siglip_forward": (C.c_int, [c_ptr, C.c_int, C.c_int, c_ptr, C.c_float, C.c_float, c_ptr]),
    "pi0_embed_language": (C.c_int, [c_ptr, C.c_int, c_ptr, c_ptr]),
    "pi0_prefill": (C.c_int, [c_ptr, C.c_int, c_ptr, c_ptr, c_ptr]),
    "pi0_denoise_loop": (C.c_int, [c_ptr, C.c_int, c_ptr, c_ptr, c_ptr]),
    "pi0_infer": (C.c_int, [c_ptr, C.c_int, C.POINTER(c_ptr), C.c_float, C.c_float, c_ptr, c_ptr, c_ptr, c_ptr,
                            c_ptr, c_ptr]),
    "pi0_debug_buffer": (c_ptr, [c_ptr, C.c_char_p, C.POINTER(C.c_int64)]),
    "pi0_profile": (C.c_int, [c_ptr, C.c_int]),
    "pi0_profile_report": (C.c_int, [c_ptr, C.c_char_p, C.c_size_t]),
    "pi0_path_report": (C.c_int, [c_ptr, C.c_int, C.c_char_p, C.c_size_t]),
    "pi0_debug_check_guards": (C.c_int, [c_ptr, C.c_char_p, C.c_size_t]),
    "pi0_launch_count": (C.c_longlong, [C.c_int]),
    "pi0_gemm": (C.c_int, [C.POINTER(GemmDesc), C.c_int, C.c_int, C.c_int, c_ptr]),
    "pi0_rmsnorm": (C.c_int, [c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, c_ptr]),
    "pi0_layernorm": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, c_ptr]),
    "pi0_softmax": (C.c_int, [c_ptr, C.c_int64, C.c_int, C.c_int64, C.c_int, C.c_int, c_ptr, C.c_int, c_ptr,
                              C.c_int, C.c_int, c_ptr]),
    "pi0_decode_attention": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_int, c_ptr]),
    "pi0_siglip_attention": (C.c_int, [c_ptr, c_ptr, C.c_int, c_ptr]),
    "pi0_prefill_attention": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.c_int, c_ptr]),
    "pi0_decode_attention_batch": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, C.c_int,
                                             C.c_int, c_ptr]),
    "pi0_gemm_f32x3": (C.c_int, [C.POINTER(GemmDesc), c_ptr, C.c_size_t, c_ptr, C.c_size_t, c_ptr, C.c_size_t, C.c_int, c_ptr]),
    "pi0_gemm_f32x3_scratch_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int]),
    "pi0_attention_long": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.c_int, C.c_int, c_ptr]),
    "pi0_decode_attention_split_static": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int,
                                                    C.c_int, C.c_int, c_ptr]),
    "pi0_decode_attention_split_static_max_bytes": (C.c_longlong, []),
    "pi0_decode_attention_split": (C.c_int, [c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, C.c_int, C.c_int, C.c_int,
                                             C.c_int, C.c_int, c_ptr]),
    "pi0_position_ids": (C.c_int, [c_ptr, C.c_int, C.c_int, C.c_int, c_ptr, c_ptr, c_ptr]),
    "pi0_adapter_resize_normalize": (C.c_int, [c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, c_ptr, c_ptr,
                                               C.c_int, c_ptr, c_ptr, C.c_int, C.c_float, C.c_float, C.c_float, c_ptr]),
    "pi0_adapter_lanczos4_normalize": (C.c_int, [c_ptr, c_ptr, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, c_ptr, c_ptr,
                                                 c_ptr, c_ptr, C.c_float, C.c_float, C.c_float, c_ptr]),
    "pi0_adapter_proprio": (C.c_int, [c_ptr, C.c_int, c_ptr, C.c_int, c_ptr, c_ptr, C.c_int, C.c_int, c_ptr]),
    "pi0_adapter_postprocess": (C.c_int, [c_ptr, c_ptr, C.c_int, c_ptr, c_ptr, C.c_int, c_ptr]),
}

_lib = None


class Pi0Error(RuntimeError):
    pass


def lib():
    """dlopen libpi0b200.so once and type every entry point. Fails loudly when the library is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Pi0Error(
            f"{LIB_PATH} not found: the CUDA extension is the only implementation of this path "
            "(no CPU fallback). Build it with `python int-act_b200/build.py`.")
    handle = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(handle, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    if handle.pi0_abi_version() != ABI_VERSION:
        raise Pi0Error(f"ABI mismatch: library {handle.pi0_abi_version()} vs binding {ABI_VERSION}")
    _lib = handle
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().pi0_last_error()
        raise Pi0Error(f"{what} failed ({rc}): {msg.decode() if msg else '?'}")


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def current_stream(device=None):
    """The calling thread's current stream ON `device` (not on the thread's current device: a server worker thread starts
    on cuda:0 whatever device the engine lives on)."""
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def gemm(A, W, out, *, M=None, N=None, K=None, lda=None, ldw=None, ldc=None, bias=None, res=None, ldr=None,
         res_row_mod=0, act=ACT_NONE, alpha=1.0, post_div=0.0, post_mul=0.0, epi_mode=EPI_STORE, bn=128,
         nz0=1, nz1=1, a_z=(0, 0), w_z=(0, 0), out_z=(0, 0), dtype=DTYPE_BF16, impl=0, num_sms=148, splits=1,
         extra_partial_bytes=0, **extra):
    """Unit-test helper around pi0_gemm: out = A . W^T with the fused epilogue (A [M,K], W [N,K] row-major)."""
    d = GemmDesc()
    d.A, d.W, d.out = ptr(A), ptr(W), ptr(out)
    d.M = M if M is not None else A.shape[-2]
    d.K = K if K is not None else A.shape[-1]
    d.N = N if N is not None else W.shape[-2]
    d.lda = lda if lda is not None else A.stride(-2)
    d.ldw = ldw if ldw is not None else W.stride(-2)
    d.ldc = ldc if ldc is not None else out.stride(-2)
    d.a_z0, d.a_z1 = a_z
    d.w_z0, d.w_z1 = w_z
    d.out_z0, d.out_z1 = out_z
    d.nz0, d.nz1 = nz0, nz1
    d.epi_mode, d.bn, d.act = epi_mode, bn, act
    d.splits = splits
    d.bias, d.res = ptr(bias), ptr(res)
    d.ldr = ldr if ldr is not None else (res.stride(-2) if res is not None else 0)
    d.res_row_mod = res_row_mod
    d.alpha, d.post_div, d.post_mul = alpha, post_div, post_mul
    d.vt_col0 = 0x7FFFFFFF
    d.vt_hd = d.vt_heads = d.vt_rpb = 1
    for k, v in extra.items():
        setattr(d, k, ptr(v) if hasattr(v, "data_ptr") else v)
    if impl == "x3":  # the fp32 regime's split-bf16 tensor-core GEMM, scratch owned by this helper
        import torch
        nz = d.nz0 * d.nz1
        ba = int(lib().pi0_gemm_f32x3_scratch_bytes(nz, d.M, d.K))
        bw = int(lib().pi0_gemm_f32x3_scratch_bytes(nz, d.N, d.K))
        sa = torch.empty(ba, dtype=torch.uint8, device=out.device)
        sw = torch.empty(bw, dtype=torch.uint8, device=out.device)
        bp = int(extra_partial_bytes)
        sp = torch.empty(max(bp, 16), dtype=torch.uint8, device=out.device)
        check(lib().pi0_gemm_f32x3(C.byref(d), ptr(sa), ba, ptr(sw), bw, ptr(sp) if bp else None, bp, num_sms,
                                   current_stream()), "pi0_gemm_f32x3")
        torch.cuda.synchronize()   # the scratch tensors die with this frame
        return out
    check(lib().pi0_gemm(C.byref(d), dtype, impl, num_sms, current_stream()), "pi0_gemm")
    return out
