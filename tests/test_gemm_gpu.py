"""GPU unit tests of the GEMM kernels through the C ABI (pi0_gemm): the tcgen05/TMEM/TMA kernel (bf16) and the
SIMT kernel (fp32 check mode / bf16 cross-check) against plain PyTorch fp32 references of the same op."""
import math

import pytest
import torch

from int_act_b200 import binding as pb

pytestmark = pytest.mark.gpu


def _ref_linear(A, W):
    return A.float() @ W.float().t()


def _rand(shape, dev, scale=1.0, dtype=torch.bfloat16, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype).to(dev)


@pytest.mark.parametrize("impl", [0, 1])
@pytest.mark.parametrize("M,N,K,bn", [(128, 128, 64, 128), (300, 512, 256, 128), (5, 1024, 2048, 32),
                                      (1000, 768, 1152, 256), (256, 72, 256, 128), (333, 2048, 4304, 64)])
def test_gemm_plain_bf16(cuda, impl, M, N, K, bn):
    A = _rand((M, K), cuda, seed=1)
    W = _rand((N, K), cuda, 0.05, seed=2)
    out = torch.zeros((M, N), dtype=torch.bfloat16, device=cuda)
    pb.gemm(A, W, out, bn=bn, impl=impl)
    torch.cuda.synchronize()
    ref = _ref_linear(A, W)
    err = (out.float() - ref).abs().max().item()
    tol = 2e-2 * ref.abs().max().item() + 1e-3
    assert err <= tol, (err, tol)


def test_gemm_tc_matches_simt_exactly_enough(cuda):
    A = _rand((700, 1024), cuda, seed=3)
    W = _rand((640, 1024), cuda, 0.03, seed=4)
    o0 = torch.zeros((700, 640), dtype=torch.bfloat16, device=cuda)
    o1 = torch.zeros_like(o0)
    pb.gemm(A, W, o0, bn=128, impl=0)
    pb.gemm(A, W, o1, bn=128, impl=1)
    torch.cuda.synchronize()
    # both round one fp32 accumulator to bf16; only the fp32 summation order differs
    assert (o0.float() - o1.float()).abs().max().item() <= 2 ** -7 * o1.float().abs().max().item()


@pytest.mark.parametrize("impl", [0, 1])
def test_gemm_bias_act_residual(cuda, impl):
    M, N, K = 520, 1152, 512
    A = _rand((M, K), cuda, seed=5)
    W = _rand((N, K), cuda, 0.05, seed=6)
    bias = _rand((N,), cuda, 0.5, torch.float32, seed=7)
    res = _rand((M, N), cuda, seed=8)
    out = torch.zeros((M, N), dtype=torch.bfloat16, device=cuda)
    pb.gemm(A, W, out, bias=bias, res=res, act=pb.ACT_GELU_TANH, bn=128, impl=impl)
    torch.cuda.synchronize()
    lin = (_ref_linear(A, W) + bias).bfloat16().float()
    ref = (torch.nn.functional.gelu(lin, approximate="tanh").bfloat16().float() + res.float()).bfloat16().float()
    assert (out.float() - ref).abs().max().item() <= 5e-2


def test_gemm_fp32_simt(cuda):
    M, N, K = 333, 520, 588
    A = _rand((M, K), cuda, dtype=torch.float32, seed=9)
    W = _rand((N, K), cuda, 0.05, torch.float32, seed=10)
    out = torch.zeros((M, N), dtype=torch.float32, device=cuda)
    pb.gemm(A, W, out, dtype=pb.DTYPE_F32)
    torch.cuda.synchronize()
    ref = (A.double() @ W.double().t()).float()
    assert (out - ref).abs().max().item() <= 1e-4


@pytest.mark.parametrize("impl", [0, 1])
def test_gemm_batched_heads(cuda, impl):
    """scores[b,h] = q[b,:,h,:] . k[b,:,h,:]^T * scale with q,k strided views of one fused qkv buffer (SigLIP)."""
    B, T, H, D = 3, 256, 16, 72
    qkv = _rand((B, T, 3 * H * D), cuda, seed=11)
    sc = torch.zeros((B, H, T, T), dtype=torch.bfloat16, device=cuda)
    q = qkv[..., : H * D]
    k = qkv[..., H * D: 2 * H * D]
    pb.gemm(q, k, sc, M=T, N=T, K=D, lda=3 * H * D, ldw=3 * H * D, ldc=T, nz0=H, nz1=B,
            a_z=(D, T * 3 * H * D), w_z=(D, T * 3 * H * D), out_z=(T * T, H * T * T), alpha=D ** -0.5, bn=128,
            impl=impl)
    torch.cuda.synchronize()
    qf = q.float().view(B, T, H, D).permute(0, 2, 1, 3)
    kf = k.float().view(B, T, H, D).permute(0, 2, 1, 3)
    ref = ((qf @ kf.transpose(2, 3)).bfloat16().float() * D ** -0.5).bfloat16().float()
    assert (sc.float() - ref).abs().max().item() <= 3e-2 * ref.abs().max().item()


@pytest.mark.parametrize("impl", [0, 1])
def test_gemm_geglu(cuda, impl):
    M, K, inter = 200, 1024, 512
    A = _rand((M, K), cuda, seed=12)
    Wg = _rand((inter, K), cuda, 0.03, seed=13)
    Wu = _rand((inter, K), cuda, 0.03, seed=14)
    # pack: per 256 rows, 128 gate rows then the matching 128 up rows
    Wgu = torch.stack([Wg.view(-1, 128, K), Wu.view(-1, 128, K)], dim=1).reshape(2 * inter, K).contiguous()
    out = torch.zeros((M, inter), dtype=torch.bfloat16, device=cuda)
    pb.gemm(A, Wgu, out, epi_mode=pb.EPI_GEGLU, bn=256, impl=impl)
    torch.cuda.synchronize()
    g = _ref_linear(A, Wg).bfloat16().float()
    u = _ref_linear(A, Wu).bfloat16().float()
    ref = (torch.nn.functional.gelu(g, approximate="tanh").bfloat16().float() * u).bfloat16().float()
    assert (out.float() - ref).abs().max().item() <= 3e-2 * max(1.0, ref.abs().max().item())


# ---- split-K (thread-block cluster + DSMEM reduction) --------------------------------------------------------------
@pytest.mark.parametrize("M,N,K,bn,splits", [(5, 1024, 4096, 64, 8), (320, 1024, 2048, 256, 8), (40, 2048, 1024, 128, 3),
                                             (328, 2048, 16384, 256, 6), (20, 1024, 1024, 32, 4), (130, 520, 1152, 256, 2),
                                             (5, 1024, 4096, 0, 0), (320, 1024, 4096, 0, 0), (256, 1152, 4304, 0, 0)])
def test_gemm_splitk_matches_unsplit(cuda, M, N, K, bn, splits):
    """The cluster split-K kernel against the fp32 reference AND against the unsplit tcgen05 kernel (same bf16 result up
    to the fp32 summation order), with bias + activation + residual in the epilogue. bn=0/splits=0 = auto pick."""
    A = _rand((M, K), cuda, seed=21)
    W = _rand((N, K), cuda, 0.03, seed=22)
    bias = _rand((N,), cuda, 0.5, torch.float32, seed=23)
    res = _rand((M, N), cuda, seed=24)
    o_split = torch.zeros((M, N), dtype=torch.bfloat16, device=cuda)
    o_plain = torch.zeros_like(o_split)
    pb.gemm(A, W, o_split, bias=bias, res=res, act=pb.ACT_SILU, bn=bn, splits=splits)
    pb.gemm(A, W, o_plain, bias=bias, res=res, act=pb.ACT_SILU, bn=bn if bn else 128, splits=1)
    torch.cuda.synchronize()
    lin = (_ref_linear(A, W) + bias).bfloat16().float()
    ref = (torch.nn.functional.silu(lin).bfloat16().float() + res.float()).bfloat16().float()
    scale = max(1.0, ref.abs().max().item())
    assert (o_split.float() - ref).abs().max().item() <= 3e-2 * scale
    assert (o_split.float() - o_plain.float()).abs().max().item() <= 2 ** -6 * scale
    # deterministic: rank-ordered DSMEM reduction
    o_again = torch.zeros_like(o_split)
    pb.gemm(A, W, o_again, bias=bias, res=res, act=pb.ACT_SILU, bn=bn, splits=splits)
    torch.cuda.synchronize()
    assert torch.equal(o_again, o_split)


@pytest.mark.parametrize("splits", [2, 4])
def test_gemm_geglu_splitk(cuda, splits):
    M, K, inter = 37, 1024, 1024
    A = _rand((M, K), cuda, seed=31)
    Wg = _rand((inter, K), cuda, 0.03, seed=32)
    Wu = _rand((inter, K), cuda, 0.03, seed=33)
    Wgu = torch.stack([Wg.view(-1, 128, K), Wu.view(-1, 128, K)], dim=1).reshape(2 * inter, K).contiguous()
    out = torch.zeros((M, inter), dtype=torch.bfloat16, device=cuda)
    pb.gemm(A, Wgu, out, epi_mode=pb.EPI_GEGLU, bn=256, splits=splits)
    torch.cuda.synchronize()
    g = _ref_linear(A, Wg).bfloat16().float()
    u = _ref_linear(A, Wu).bfloat16().float()
    ref = (torch.nn.functional.gelu(g, approximate="tanh").bfloat16().float() * u).bfloat16().float()
    assert (out.float() - ref).abs().max().item() <= 3e-2 * max(1.0, ref.abs().max().item())


def test_gemm_batched_splitk(cuda):
    """Denoise-shaped scores: 40 query rows x 333 keys per sample, batched over samples, K = 256 split in two."""
    B, R, Kk, D = 6, 40, 333, 256
    q = _rand((B, R, D), cuda, seed=41)
    k = _rand((B, 336, D), cuda, seed=42)
    sc = torch.zeros((B, R, 336), dtype=torch.bfloat16, device=cuda)
    pb.gemm(q, k, sc, M=R, N=Kk, K=D, lda=D, ldw=D, ldc=336, nz0=B, a_z=(R * D, 0), w_z=(336 * D, 0),
            out_z=(R * 336, 0), alpha=0.0625, bn=64, splits=2)
    torch.cuda.synchronize()
    ref = ((q.float() @ k[:, :Kk].float().transpose(1, 2)).bfloat16().float() * 0.0625).bfloat16().float()
    assert (sc[:, :, :Kk].float() - ref).abs().max().item() <= 3e-2 * ref.abs().max().item()


# ---- CTA-pair kernel (tcgen05 cta_group::2): impl 3 forces it, impl 2 forbids it -----------------------------------------
@pytest.mark.parametrize("M,N,K,bn", [(256, 256, 64, 256), (1000, 768, 1152, 256), (2304, 4304, 1152, 256),
                                      (700, 1152, 4304, 128), (130, 520, 192, 256), (4096, 2048, 2048, 128)])
def test_gemm_pair_matches_single_cta(cuda, M, N, K, bn):
    """Same tile arithmetic on one SM or on a CTA pair: bias + activation + residual epilogue, ragged M / N / K tails,
    several tiles per pair (persistent loop, both TMEM accumulator buffers, ring wrap-around)."""
    A = _rand((M, K), cuda, seed=51)
    W = _rand((N, K), cuda, 0.03, seed=52)
    bias = _rand((N,), cuda, 0.5, torch.float32, seed=53)
    res = _rand((M, N), cuda, seed=54)
    o_pair = torch.zeros((M, N), dtype=torch.bfloat16, device=cuda)
    o_one = torch.zeros_like(o_pair)
    pb.gemm(A, W, o_pair, bias=bias, res=res, act=pb.ACT_GELU_TANH, bn=bn, impl=3)
    pb.gemm(A, W, o_one, bias=bias, res=res, act=pb.ACT_GELU_TANH, bn=bn, impl=2)
    torch.cuda.synchronize()
    lin = (_ref_linear(A, W) + bias).bfloat16().float()
    ref = (torch.nn.functional.gelu(lin, approximate="tanh").bfloat16().float() + res.float()).bfloat16().float()
    scale = max(1.0, ref.abs().max().item())
    assert (o_pair.float() - ref).abs().max().item() <= 3e-2 * scale
    assert (o_pair.float() - o_one.float()).abs().max().item() <= 2 ** -6 * scale


def test_gemm_pair_few_sms(cuda):
    """A grid of 2 pairs walks 40 tiles: exercises the persistent tile loop and accumulator hand-back far past one wave."""
    M, N, K = 1280, 2048, 512
    A = _rand((M, K), cuda, seed=55)
    W = _rand((N, K), cuda, 0.03, seed=56)
    out = torch.zeros((M, N), dtype=torch.bfloat16, device=cuda)
    pb.gemm(A, W, out, bn=256, impl=3, num_sms=4)
    torch.cuda.synchronize()
    ref = _ref_linear(A, W)
    assert (out.float() - ref).abs().max().item() <= 2e-2 * ref.abs().max().item() + 1e-3


def test_gemm_pair_geglu(cuda):
    M, K, inter = 1500, 1024, 2048
    A = _rand((M, K), cuda, seed=57)
    Wg = _rand((inter, K), cuda, 0.03, seed=58)
    Wu = _rand((inter, K), cuda, 0.03, seed=59)
    Wgu = torch.stack([Wg.view(-1, 128, K), Wu.view(-1, 128, K)], dim=1).reshape(2 * inter, K).contiguous()
    out = torch.zeros((M, inter), dtype=torch.bfloat16, device=cuda)
    one = torch.zeros_like(out)
    pb.gemm(A, Wgu, out, epi_mode=pb.EPI_GEGLU, bn=256, impl=3)
    pb.gemm(A, Wgu, one, epi_mode=pb.EPI_GEGLU, bn=256, impl=2)
    torch.cuda.synchronize()
    g = _ref_linear(A, Wg).bfloat16().float()
    u = _ref_linear(A, Wu).bfloat16().float()
    ref = (torch.nn.functional.gelu(g, approximate="tanh").bfloat16().float() * u).bfloat16().float()
    scale = max(1.0, ref.abs().max().item())
    assert (out.float() - ref).abs().max().item() <= 3e-2 * scale
    assert (out.float() - one.float()).abs().max().item() <= 2 ** -6 * scale


def test_gemm_pair_rope_qkv(cuda):
    """Fused q|k|v projection with the RoPE / KV-cache epilogue: the pair kernel writes the same q, K cache and V^T cache
    as the single-CTA kernel (which the end-to-end tests pin against the oracle)."""
    B, T, hid, heads = 3, 328, 512, 8
    M, N = B * T, (heads + 2) * 256
    A = _rand((M, hid), cuda, seed=60)
    W = _rand((N, hid), cuda, 0.05, seed=61)
    pos = (torch.arange(M, device=cuda, dtype=torch.int32) % T).contiguous()
    inv = 1.0 / (10000.0 ** (torch.arange(0, 128, device=cuda, dtype=torch.float32) / 128.0))
    ang = torch.arange(T, device=cuda, dtype=torch.float32)[:, None] * inv[None, :]
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    kv_pitch, vt_pitch = 336, 336
    outs = []
    for impl in (3, 2):
        q = torch.zeros((M, heads * 256), dtype=torch.bfloat16, device=cuda)
        kc = torch.zeros((B, kv_pitch, 256), dtype=torch.bfloat16, device=cuda)
        vt = torch.zeros((B, 256, vt_pitch), dtype=torch.bfloat16, device=cuda)
        pb.gemm(A, W, q, epi_mode=pb.EPI_ROPE, bn=256, impl=impl, ldc=heads * 256, rope_cos=cos, rope_sin=sin,
                pos_ids=pos, kcache=kc, n_q_heads=heads, kv_pitch=kv_pitch, vt=vt, vt_pitch=vt_pitch, vt_rpb=T, vt_toff=0)
        torch.cuda.synchronize()
        outs.append((q, kc, vt))
    lin = _ref_linear(A, W).bfloat16().float().view(B, T, heads + 2, 256)
    c, s_ = cos[None, :, None, :], sin[None, :, None, :]
    x1, x2 = lin[..., :128], lin[..., 128:]
    rot = torch.cat([x1 * c - x2 * s_, x2 * c + x1 * s_], dim=-1).bfloat16().float()
    assert (outs[0][0].float().view(B, T, heads, 256) - rot[:, :, :heads]).abs().max().item() <= 5e-2
    assert (outs[0][1][:, :T].float() - rot[:, :, heads]).abs().max().item() <= 5e-2
    assert (outs[0][2][:, :, :T].float() - lin[:, :, heads + 1].transpose(1, 2)).abs().max().item() <= 5e-2
    for a, b in zip(outs[0], outs[1]):
        assert (a.float() - b.float()).abs().max().item() <= 2 ** -6 * max(1.0, b.float().abs().max().item())


def test_gemm_tc_throughput_report(cuda):
    """Not an assertion on speed: prints achieved TFLOP/s of the tcgen05 kernel for the log."""
    M, N, K = 8192, 8192, 2048
    A = _rand((M, K), cuda, seed=15)
    W = _rand((N, K), cuda, 0.02, seed=16)
    out = torch.zeros((M, N), dtype=torch.bfloat16, device=cuda)
    for bn in (128, 256):
        for _ in range(3):
            pb.gemm(A, W, out, bn=bn)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            pb.gemm(A, W, out, bn=bn)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"\n[tc gemm] M={M} N={N} K={K} bn={bn}: {ms:.3f} ms  {2.0 * M * N * K / ms / 1e9:.1f} TFLOP/s")
    ref = _ref_linear(A[:256], W[:256])
    assert (out[:256, :256].float() - ref).abs().max().item() <= 2e-2 * ref.abs().max().item() + 1e-3


# ---- GemmaRMSNorm fused into the A operand (rollout-batch denoise path) -------------------------------------------------
@pytest.mark.parametrize("M,K,inter,splits", [(320, 1024, 1024, 1), (40, 1024, 512, 2), (130, 2048, 256, 4)])
def test_gemm_fused_rmsnorm_matches_norm_then_gemm(cuda, M, K, inter, splits):
    """Producer: a residual GEMM whose epilogue leaves per-chunk sums of squares of its output rows (rowsq_out). Consumer: the
    gate/up GeGLU GEMM with the RAW stream as A and the RMSNorm applied in shared memory (norm_gain / norm_rowsq), against
    pi0_rmsnorm followed by the plain GEMM - same arithmetic, so the two agree up to the summation order of the row statistic."""
    Kp = 512
    A0 = _rand((M, Kp), cuda, seed=31)
    Wp = _rand((K, Kp), cuda, 0.04, seed=32)
    res = _rand((M, K), cuda, seed=33)
    y = torch.zeros((M, K), dtype=torch.bfloat16, device=cuda)
    rowsq = torch.full((M, K // 32), float("nan"), dtype=torch.float32, device=cuda)
    pb.gemm(A0, Wp, y, res=res, bn=0, splits=0, rowsq_out=rowsq, rowsq_ld=K // 32)
    torch.cuda.synchronize()
    ref_sq = (y.float() ** 2).view(M, K // 32, 32).sum(-1)
    assert torch.isfinite(rowsq).all()
    assert (rowsq - ref_sq).abs().max().item() <= 1e-3 * ref_sq.abs().max().item()

    gain = (1.0 + 0.1 * torch.randn(K, generator=torch.Generator().manual_seed(34))).to(cuda)
    Wg = _rand((inter, K), cuda, 0.03, seed=35)
    Wu = _rand((inter, K), cuda, 0.03, seed=36)
    Wgu = torch.stack([Wg.view(-1, 128, K), Wu.view(-1, 128, K)], dim=1).reshape(2 * inter, K).contiguous()
    fused = torch.zeros((M, inter), dtype=torch.bfloat16, device=cuda)
    pb.gemm(y, Wgu, fused, epi_mode=pb.EPI_GEGLU, bn=256, splits=splits, norm_gain=gain, norm_rowsq=rowsq,
            norm_slots=K // 32, norm_eps=1e-6)
    h = torch.zeros_like(y)
    pb.check(pb.lib().pi0_rmsnorm(pb.ptr(y), pb.ptr(h), pb.ptr(gain), M, K, 0, pb.current_stream()), "pi0_rmsnorm")
    plain = torch.zeros_like(fused)
    pb.gemm(h, Wgu, plain, epi_mode=pb.EPI_GEGLU, bn=256, splits=splits)
    torch.cuda.synchronize()
    assert torch.isfinite(fused.float()).all()
    dif = (fused.float() - plain.float()).abs().max().item()
    scale = max(1.0, plain.float().abs().max().item())
    assert dif <= 2 ** -7 * scale, dif          # at most a flipped bf16 rounding here and there
    assert (fused != plain).float().mean().item() < 0.02
    hh = (y.float() * torch.rsqrt((y.float() ** 2).mean(-1, keepdim=True) + 1e-6) * gain).bfloat16().float()
    g = (hh @ Wg.float().t()).bfloat16().float()
    u = (hh @ Wu.float().t()).bfloat16().float()
    ref = (torch.nn.functional.gelu(g, approximate="tanh").bfloat16().float() * u).bfloat16().float()
    assert (fused.float() - ref).abs().max().item() <= 3e-2 * max(1.0, ref.abs().max().item())
