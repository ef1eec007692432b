"""The GEMM of the fp32 regime on the tensor cores (csrc/gemm_x3.cu + gemm_tc_seg_kernel, through pi0_gemm_f32x3): fp32 operands
split into three bf16 pieces each, six piece products accumulated in fp32 with segmented accumulation. Checked against fp64
products at the tolerance of the fp32 regime (<= 1e-4 on O(1)-O(10) values; measured ~1e-6 relative) and against the SIMT check
kernel with the same fused epilogues."""
import pytest
import torch

from int_act_b200 import binding as pb

pytestmark = pytest.mark.gpu


def _rand(shape, dev, scale=1.0, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dev)


@pytest.mark.parametrize("M,N,K", [(333, 520, 588), (5, 1024, 2048), (1000, 768, 1152), (328, 2048, 16384), (130, 72, 250),
                                   (257, 32, 4304)])
def test_x3_plain_matches_fp64(cuda, M, N, K):
    A = _rand((M, K), cuda, seed=1)
    W = _rand((N, K), cuda, K ** -0.5, seed=2)
    out = torch.full((M, N), float("nan"), device=cuda)
    pb.gemm(A, W, out, dtype=pb.DTYPE_F32, impl="x3")
    ref = A.double() @ W.double().t()
    err = (out.double() - ref).abs().max().item()
    t32 = (A @ W.t()).double()
    print(f"\n[x3 gemm M={M} N={N} K={K}] max err vs fp64 {err:.2e} (torch fp32 matmul: {(t32 - ref).abs().max().item():.2e})")
    assert err <= 2e-5, err


@pytest.mark.parametrize("M,N,K", [(40, 1024, 4096), (5, 2560, 1024), (328, 2048, 2048), (16, 520, 16384)])
def test_x3_split_k_over_idle_sms_matches_fp64(cuda, M, N, K):
    """Skinny problems (denoise steps, batch-1 prefill): with a partial-sum workspace every CTA takes a k-slice and a second
    launch sums the slices in split order and runs the epilogue (bias + residual here)."""
    A = _rand((M, K), cuda, seed=21)
    W = _rand((N, K), cuda, K ** -0.5, seed=22)
    bias = _rand((N,), cuda, 0.1, seed=23)
    res = _rand((M, N), cuda, seed=24)
    outs = []
    for pbytes in (64 << 20, 0):
        out = torch.full((M, N), float("nan"), device=cuda)
        pb.gemm(A, W, out, bias=bias, res=res, dtype=pb.DTYPE_F32, impl="x3", bn=256, extra_partial_bytes=pbytes)
        outs.append(out)
    ref = A.double() @ W.double().t() + bias.double() + res.double()
    assert (outs[0].double() - ref).abs().max().item() <= 2e-5
    assert (outs[0] - outs[1]).abs().max().item() <= 1e-5      # split and unsplit forms agree to fp32 summation order
    again = torch.zeros((M, N), device=cuda)
    pb.gemm(A, W, again, bias=bias, res=res, dtype=pb.DTYPE_F32, impl="x3", bn=256, extra_partial_bytes=64 << 20)
    assert torch.equal(again, outs[0])                          # deterministic


def test_x3_geglu_split_k(cuda):
    M, K, inter = 40, 1024, 1024
    A = _rand((M, K), cuda, seed=12)
    Wg = _rand((inter, K), cuda, 0.03, seed=13)
    Wu = _rand((inter, K), cuda, 0.03, seed=14)
    Wgu = torch.stack([Wg.view(-1, 128, K), Wu.view(-1, 128, K)], dim=1).reshape(2 * inter, K).contiguous()
    out = torch.zeros((M, inter), device=cuda)
    pb.gemm(A, Wgu, out, epi_mode=pb.EPI_GEGLU, bn=256, dtype=pb.DTYPE_F32, impl="x3", extra_partial_bytes=64 << 20)
    ref = torch.nn.functional.gelu(A.double() @ Wg.double().t(), approximate="tanh") * (A.double() @ Wu.double().t())
    assert (out.double() - ref).abs().max().item() <= 1e-4


def test_x3_split_is_exact_for_fp32_inputs_with_24_significant_bits(cuda):
    """hi + mid + lo reproduces every fp32 value: with a one-hot W the GEMM must return A bit for bit (every dropped piece
    product is zero then)."""
    M, K = 200, 256
    A = _rand((M, K), cuda, 3.0, seed=3)
    W = torch.eye(K, device=cuda)
    out = torch.zeros((M, K), device=cuda)
    pb.gemm(A, W, out, dtype=pb.DTYPE_F32, impl="x3")
    assert torch.equal(out, A)


def test_x3_bias_act_residual_matches_simt(cuda):
    M, N, K = 300, 384, 1152
    A = _rand((M, K), cuda, seed=4)
    W = _rand((N, K), cuda, 0.03, seed=5)
    bias = _rand((N,), cuda, 0.1, seed=6)
    res = _rand((M, N), cuda, seed=7)
    outs = []
    for impl in ("x3", 0):
        out = torch.zeros((M, N), device=cuda)
        pb.gemm(A, W, out, bias=bias, res=res, act=pb.ACT_GELU_TANH, dtype=pb.DTYPE_F32, impl=impl)
        torch.cuda.synchronize()
        outs.append(out)
    ref = torch.nn.functional.gelu((A.double() @ W.double().t()) + bias.double(), approximate="tanh") + res.double()
    assert (outs[0].double() - ref).abs().max().item() <= 1e-4   # fp32 tanh of the shared epilogue against an fp64 one
    assert (outs[0] - outs[1]).abs().max().item() <= 5e-6 * max(1.0, ref.abs().max().item())


def test_x3_batched_heads_with_strided_operands(cuda):
    """scores[b, h] = q . k^T * scale with q, k strided views of one fused qkv buffer (SigLIP shapes, K = 72)."""
    B, T, H, D = 2, 256, 16, 72
    qkv = _rand((B, T, 3 * H * D), cuda, seed=11)
    sc = torch.zeros((B, H, T, T), device=cuda)
    q = qkv[..., : H * D]
    k = qkv[..., H * D: 2 * H * D]
    pb.gemm(q, k, sc, M=T, N=T, K=D, lda=3 * H * D, ldw=3 * H * D, ldc=T, nz0=H, nz1=B,
            a_z=(D, T * 3 * H * D), w_z=(D, T * 3 * H * D), out_z=(T * T, H * T * T), alpha=D ** -0.5, bn=128,
            dtype=pb.DTYPE_F32, impl="x3")
    qd = q.double().view(B, T, H, D).permute(0, 2, 1, 3)
    kd = k.double().view(B, T, H, D).permute(0, 2, 1, 3)
    ref = (qd @ kd.transpose(2, 3)) * D ** -0.5
    assert (sc.double() - ref).abs().max().item() <= 2e-5


def test_x3_geglu_matches_simt(cuda):
    M, K, inter = 200, 1024, 512
    A = _rand((M, K), cuda, seed=12)
    Wg = _rand((inter, K), cuda, 0.03, seed=13)
    Wu = _rand((inter, K), cuda, 0.03, seed=14)
    Wgu = torch.stack([Wg.view(-1, 128, K), Wu.view(-1, 128, K)], dim=1).reshape(2 * inter, K).contiguous()
    outs = []
    for impl in ("x3", 0):
        out = torch.zeros((M, inter), device=cuda)
        pb.gemm(A, Wgu, out, epi_mode=pb.EPI_GEGLU, bn=256, dtype=pb.DTYPE_F32, impl=impl)
        torch.cuda.synchronize()
        outs.append(out)
    ref = torch.nn.functional.gelu(A.double() @ Wg.double().t(), approximate="tanh") * (A.double() @ Wu.double().t())
    assert (outs[0].double() - ref).abs().max().item() <= 1e-4   # fp32 tanh of the shared epilogue against an fp64 one
    assert (outs[0] - outs[1]).abs().max().item() <= 5e-6 * max(1.0, ref.abs().max().item())


def test_x3_rejects_a_problem_larger_than_its_scratch(cuda):
    import ctypes as C
    A = _rand((64, 64), cuda, seed=1)
    out = torch.zeros((64, 64), device=cuda)
    d = pb.GemmDesc()
    d.A, d.W, d.out = pb.ptr(A), pb.ptr(A), pb.ptr(out)
    d.M = d.N = d.K = d.lda = d.ldw = d.ldc = 64
    d.nz0 = d.nz1 = 1
    d.bn, d.alpha = 256, 1.0
    d.vt_col0 = 0x7FFFFFFF
    d.vt_hd = d.vt_heads = d.vt_rpb = 1
    small = torch.empty(1024, dtype=torch.uint8, device=cuda)
    rc = pb.lib().pi0_gemm_f32x3(C.byref(d), pb.ptr(small), 1024, pb.ptr(small), 1024, None, 0, 148, pb.current_stream())
    assert rc == -2 and "scratch" in pb.lib().pi0_last_error().decode()
