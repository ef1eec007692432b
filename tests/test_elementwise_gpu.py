"""GPU unit tests of the HBM-bound kernels through the C ABI (pi0_rmsnorm / pi0_layernorm / pi0_softmax /
pi0_position_ids) against plain PyTorch fp32 references of the same op."""
import ctypes as C

import pytest
import torch

from int_act_b200 import binding as pb

pytestmark = pytest.mark.gpu
NEG = -2.3819763e38


def _stream():
    return pb.current_stream()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("M,D", [(5, 1024), (333, 2048), (17, 4096), (9, 72)])
def test_rmsnorm(cuda, dtype, M, D):
    g = torch.Generator().manual_seed(1)
    x = torch.randn((M, D), generator=g).to(dtype).to(cuda)
    gain = (1.0 + 0.1 * torch.randn(D, generator=g)).to(cuda)
    y = torch.empty_like(x)
    pb.check(pb.lib().pi0_rmsnorm(pb.ptr(x), pb.ptr(y), pb.ptr(gain), M, D, int(dtype == torch.float32), _stream()))
    torch.cuda.synchronize()
    xf = x.float()
    ref = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-6) * gain).to(dtype).float()
    tol = 1e-5 if dtype == torch.float32 else 2 ** -7
    assert (y.float() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("M,D", [(256, 1152), (13, 1152), (7, 2048)])
def test_layernorm(cuda, dtype, M, D):
    g = torch.Generator().manual_seed(2)
    x = (torch.randn((M, D), generator=g) * 2 + 0.3).to(dtype).to(cuda)
    w = (1.0 + 0.1 * torch.randn(D, generator=g)).to(cuda)
    b = (0.1 * torch.randn(D, generator=g)).to(cuda)
    y = torch.empty_like(x)
    pb.check(pb.lib().pi0_layernorm(pb.ptr(x), pb.ptr(y), pb.ptr(w), pb.ptr(b), M, D, int(dtype == torch.float32),
                                    _stream()))
    torch.cuda.synchronize()
    ref = torch.nn.functional.layer_norm(x.float(), (D,), w, b, 1e-6).to(dtype).float()
    tol = 2e-5 if dtype == torch.float32 else 2 ** -7
    assert (y.float() - ref).abs().max().item() <= tol * max(1.0, ref.abs().max().item())


def _softmax_ref(sc, pad, cum, n_prefix, heads, dtype):
    """sc [B, R, n_keys] fp32 (already rounded to dtype); pad [B, n_prefix] bool or None; cum [S] or None."""
    B, R, n_keys = sc.shape
    mask = torch.ones((B, R, n_keys), dtype=torch.bool)
    if pad is not None:
        mask[:, :, :n_prefix] = pad[:, None, :]
    if cum is not None:
        q = cum[torch.arange(R) // heads]                      # [R]
        mask[:, :, n_prefix:] = (cum[None, :] <= q[:, None])[None]
    x = torch.where(mask, sc, torch.tensor(NEG))
    return torch.softmax(x, dim=-1).to(dtype).float()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("B,T,P,S", [(3, 328, 328, 0), (2, 5, 328, 5), (2, 17, 584, 17), (2, 256, 256, 0),
                                     (1, 5, 1096, 5)])
def test_masked_softmax(cuda, dtype, B, T, P, S):
    """Prefix rows (pad gate only), suffix rows (pad gate + block-causal cumsum rule), SigLIP rows (no mask), and a
    row longer than the register-resident kernel covers (generic fallback)."""
    heads = 8
    g = torch.Generator().manual_seed(3)
    n_keys = P + S
    pitch = (n_keys + 7) // 8 * 8
    R = T * heads
    sc = torch.zeros((B, R, pitch), dtype=dtype)
    sc[:, :, :n_keys] = (torch.randn((B, R, n_keys), generator=g) * 3).to(dtype)
    no_mask = P == 256
    pad = None
    if not no_mask:
        pad = torch.ones((B, P), dtype=torch.bool)
        pad[0, P - 60:] = False        # right-padded text
        if B > 1:
            pad[1, P - 1:] = False
    cum = torch.tensor([1] + [2] * (S - 1), dtype=torch.int32) if S > 0 else None
    d_sc = sc.to(cuda)
    d_pad = pad.to(torch.uint8).to(cuda) if pad is not None else None
    d_cum = cum.to(cuda) if cum is not None else None
    pb.check(pb.lib().pi0_softmax(pb.ptr(d_sc), B * R, R, pitch, n_keys, P, pb.ptr(d_pad), P, pb.ptr(d_cum),
                                  heads if S > 0 else 1, int(dtype == torch.float32), _stream()))
    torch.cuda.synchronize()
    ref = _softmax_ref(sc[:, :, :n_keys].float(), pad, cum, P, heads, dtype)
    got = d_sc.cpu().float()[:, :, :n_keys]
    tol = 1e-6 if dtype == torch.float32 else 2 ** -8
    assert (got - ref).abs().max().item() <= tol
    # masked keys carry exactly zero probability, rows sum to one
    if pad is not None:
        assert got[0, :, P - 60: P].abs().max().item() == 0.0
    assert (got.sum(-1) - 1).abs().max().item() <= (1e-5 if dtype == torch.float32 else 2e-2)


def test_fully_masked_row_is_uniform(cuda):
    """The reference masks with a finite value, so a row whose keys are all masked is uniform, not NaN
    ([mirror] mllm.py:301-307)."""
    P = 64
    sc = torch.randn((1, 8, P)).bfloat16().to(cuda)
    pad = torch.zeros((1, P), dtype=torch.uint8, device=cuda)
    pb.check(pb.lib().pi0_softmax(pb.ptr(sc), 8, 8, P, P, P, pb.ptr(pad), P, None, 1, 0, _stream()))
    torch.cuda.synchronize()
    assert torch.allclose(sc.float().cpu(), torch.full((1, 8, P), 1.0 / P).bfloat16().float())


def test_position_ids_bit_exact(cuda):
    B, P, S = 4, 328, 5
    g = torch.Generator().manual_seed(4)
    pad = torch.rand((B, P), generator=g) > 0.2
    pad[3, :10] = False  # leading pads give -1
    d_pad = pad.to(torch.uint8).to(cuda)
    pp = torch.empty((B, P), dtype=torch.int32, device=cuda)
    sp = torch.empty((B, S), dtype=torch.int32, device=cuda)
    pb.check(pb.lib().pi0_position_ids(pb.ptr(d_pad), B, P, S, pb.ptr(pp), pb.ptr(sp), _stream()))
    torch.cuda.synchronize()
    assert torch.equal(pp.cpu().long(), torch.cumsum(pad.long(), 1) - 1)
    assert torch.equal(sp.cpu().long(), pad.sum(1)[:, None] + torch.arange(S)[None])


@pytest.mark.parametrize("B,S,P", [(1, 5, 328), (3, 5, 328), (70, 5, 320), (2, 17, 584), (2, 1, 64)])
def test_decode_attention_fused(cuda, B, S, P):
    """Fused suffix attention (scores + masked softmax + P.V in one launch) against a torch restatement with the
    reference's rounding points ([mirror] mllm.py:265-317): bf16 scores, fp32 softmax, bf16 probabilities."""
    g = torch.Generator().manual_seed(5)
    n_keys = P + S
    kv_pitch, vt_pitch = n_keys, (n_keys + 7) // 8 * 8
    q = torch.randn((B * S, 8 * 256), generator=g).bfloat16()
    k = torch.randn((B, kv_pitch, 256), generator=g).bfloat16()
    v = torch.randn((B, n_keys, 256), generator=g).bfloat16()
    vt = torch.zeros((B, 256, vt_pitch), dtype=torch.bfloat16)
    vt[:, :, :n_keys] = v.transpose(1, 2)
    pad = torch.ones((B, P), dtype=torch.bool)
    pad[0, P - 37:] = False
    if B > 1:
        pad[1, :3] = False
    cum = torch.tensor([1] + [2] * (S - 1), dtype=torch.int32)
    d = lambda t: t.to(cuda)
    att = torch.zeros((B * S, 8 * 256), dtype=torch.bfloat16, device=cuda)
    dq, dk, dvt, dpad, dcum = d(q), d(k), d(vt), d(pad.to(torch.uint8)), d(cum)
    pb.check(pb.lib().pi0_decode_attention(pb.ptr(dq), pb.ptr(dk), pb.ptr(dvt), pb.ptr(att), pb.ptr(dpad), pb.ptr(dcum),
                                           B, S, P, kv_pitch, vt_pitch, 148, _stream()), "pi0_decode_attention")
    torch.cuda.synchronize()
    qf = q.float().view(B, S, 8, 256)
    sc = torch.einsum("bshd,bjd->bshj", qf, k.float()).bfloat16().float()
    sc = (sc * 0.0625).bfloat16().float()
    mask = torch.ones((B, S, n_keys), dtype=torch.bool)
    mask[:, :, :P] = pad[:, None, :]
    mask[:, :, P:] = (cum[None, :] <= cum[:, None])[None]
    sc = torch.where(mask[:, :, None, :], sc, torch.tensor(NEG))
    pr = torch.softmax(sc, dim=-1).bfloat16().float()
    ref = torch.einsum("bshj,bjd->bshd", pr, v.float()).bfloat16().float().reshape(B * S, 8 * 256)
    err = (att.float().cpu() - ref).abs().max().item()
    assert err <= 2 ** -6 * max(1.0, ref.abs().max().item()), err
    if S <= 6 and n_keys <= 384:
        # the two-launch split form (rollout batches) must agree with the fused kernel (same arithmetic, same order
        # inside every dot product up to the mma k-grouping)
        att2 = torch.zeros_like(att)
        scratch = torch.zeros((B * 8 * S, (n_keys + 31) // 32 * 32), dtype=torch.bfloat16, device=cuda)
        pb.check(pb.lib().pi0_decode_attention_split(pb.ptr(dq), pb.ptr(dk), pb.ptr(dvt), pb.ptr(att2), pb.ptr(scratch),
                                                     pb.ptr(dpad), pb.ptr(dcum), B, S, P, kv_pitch, vt_pitch, _stream()),
                 "pi0_decode_attention_split")
        torch.cuda.synchronize()
        err2 = (att2.float().cpu() - ref).abs().max().item()
        assert err2 <= 2 ** -6 * max(1.0, ref.abs().max().item()), err2
    if S <= 6 and n_keys <= 352:
        # one CTA per sample, K / V^T staged in shared memory (the rollout-batch path)
        att3 = torch.zeros_like(att)
        pb.check(pb.lib().pi0_decode_attention_batch(pb.ptr(dq), pb.ptr(dk), pb.ptr(dvt), pb.ptr(att3), pb.ptr(dpad),
                                                     pb.ptr(dcum), B, S, P, kv_pitch, vt_pitch, _stream()),
                 "pi0_decode_attention_batch")
        torch.cuda.synchronize()
        err3 = (att3.float().cpu() - ref).abs().max().item()
        assert err3 <= 2 ** -6 * max(1.0, ref.abs().max().item()), err3


@pytest.mark.parametrize("mode", ["tc", "mma"])
@pytest.mark.parametrize("B", [1, 3, 12])
def test_siglip_attention_fused(cuda, B, mode, monkeypatch):
    """Fused SigLIP attention against a torch restatement with the reference's rounding points: bf16 scores scaled by
    72^-0.5, fp32 softmax, bf16 probabilities, bf16 output. mode "tc" = the tcgen05 kernel (attn_siglip_tc.cu, the product
    path; B = 12 gives every CTA of the persistent grid more than one (image, head) item), "mma" = the mma.sync kernel it
    replaced (attn_siglip.cu, PI0_SIGLIP_ATTN=mma)."""
    monkeypatch.setenv("PI0_SIGLIP_ATTN", mode)
    g = torch.Generator().manual_seed(9)
    T, H, D = 256, 16, 72
    qk = (torch.randn((B * T, 2 * H * D), generator=g) * 1.5).bfloat16()
    v = torch.randn((B, H, T, D), generator=g).bfloat16()
    qkv = torch.cat([qk, v.transpose(1, 2).reshape(B * T, H * D)], dim=1).contiguous()     # [B*256, 3456]
    att = torch.full((B * T, H * D), float("nan"), dtype=torch.bfloat16, device=cuda)
    dqkv = qkv.to(cuda)
    for _ in range(2):   # the second launch re-uses the TMEM / barrier state conventions from a fresh kernel: must agree
        pb.check(pb.lib().pi0_siglip_attention(pb.ptr(dqkv), pb.ptr(att), B, _stream()), "pi0_siglip_attention")
    torch.cuda.synchronize()
    q = qk[:, : H * D].float().view(B, T, H, D).transpose(1, 2)
    k = qk[:, H * D:].float().view(B, T, H, D).transpose(1, 2)
    sc = (q @ k.transpose(2, 3)).bfloat16().float()
    sc = (sc * torch.tensor(72 ** -0.5, dtype=torch.float32)).bfloat16().float()
    pr = torch.softmax(sc, dim=-1).bfloat16().float()
    ref = (pr @ v.float()).bfloat16().float().transpose(1, 2).reshape(B * T, H * D)
    assert torch.isfinite(att.float()).all()
    err = (att.float().cpu() - ref).abs().max().item()
    assert err <= 2 ** -6 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("B,T,P,pitch,sms", [(1, 328, 328, 336, 148), (3, 328, 328, 333 + 3, 148), (2, 72, 72, 80, 148),
                                             (2, 130, 136, 144, 148), (3, 328, 328, 336, 5), (5, 200, 208, 216, 3)])
def test_prefill_attention_fused(cuda, B, T, P, pitch, sms):
    """Fused tcgen05 prefix attention (attn_prefill.cu) against a torch restatement with the reference's rounding
    points: bf16 scores scaled by 256^-0.5, key mask with the finite mask value, fp32 softmax, bf16 probabilities, bf16
    output. MQA: 8 query heads per token against one K / V; ragged tiles (T * 8 not a multiple of 128), padded keys,
    stale suffix rows behind the prefix keys in both caches, a fully padded key range and rows whose token is itself pad.
    T <= P: the queries are the first T prefix tokens. sms < tiles: the persistent loop (TMEM / barrier phases, staging
    buffers aliased with the P tile) over many tiles per CTA, across sample boundaries."""
    g = torch.Generator().manual_seed(11)
    q = (torch.randn((B * T, 8 * 256), generator=g) * 0.8).bfloat16()
    kc = torch.randn((B, pitch, 256), generator=g).bfloat16()           # rows >= P: stale data that must not be read
    vt = torch.randn((B, 256, pitch), generator=g).bfloat16()
    pad = torch.ones((B, P), dtype=torch.uint8)
    pad[0, P - 17:] = 0                                                 # right-padded language tokens
    if B > 1:
        pad[1, 5:40] = 0
    att = torch.full((B * T, 8 * 256), 7.0, dtype=torch.bfloat16, device=cuda)
    dq, dk, dv, dp = q.to(cuda), kc.to(cuda), vt.to(cuda), pad.to(cuda)
    pb.check(pb.lib().pi0_prefill_attention(pb.ptr(dq), pb.ptr(dk), pb.ptr(dv), pb.ptr(att), pb.ptr(dp), B, T, P, pitch,
                                            pitch, sms, _stream()), "pi0_prefill_attention")
    torch.cuda.synchronize()
    qf = q.float().view(B, T * 8, 256)
    sc = (qf @ kc[:, :P].float().transpose(1, 2)).bfloat16().float()
    sc = (sc * 0.0625).bfloat16().float()
    sc = torch.where(pad[:, None, :].bool(), sc, torch.tensor(-2.3819763e38))
    pr = torch.softmax(sc, dim=-1).bfloat16().float()
    ref = (pr @ vt[:, :, :P].float().transpose(1, 2)).bfloat16().float().view(B * T, 8 * 256)
    err = (att.float().cpu() - ref).abs().max().item()
    assert err <= 2 ** -6 * max(1.0, ref.abs().max().item()), err
