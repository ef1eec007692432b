"""The long-key tcgen05 attention (csrc/attn_long.cu, through pi0_attention_long) against torch restatements with the
reference's rounding points (bf16 scores scaled by 256^-0.5, finite mask value, fp32 softmax, bf16 probabilities, bf16
output): prefix mode at the LIBERO (P = 584) and 3-camera (P = 840) shapes, suffix mode (17 suffix tokens over P + 17 keys,
pad mask + cumsum rule) also against the existing one-CTA-per-sample kernel (pi0_decode_attention)."""
import pytest
import torch

from int_act_b200 import binding as pb

pytestmark = pytest.mark.gpu
NEG = -2.3819763e38


def _stream():
    return pb.current_stream()


@pytest.mark.parametrize("B,T,P,pitch,sms", [(2, 584, 584, 608, 148), (1, 840, 840, 864, 148), (3, 72, 328, 336, 148),
                                             (3, 584, 584, 601 + 7, 7), (2, 100, 1000, 1024, 148)])
def test_prefix_mode(cuda, B, T, P, pitch, sms):
    """T <= P: the queries are the first T prefix tokens. Ragged last tile (T * 8 % 128 != 0), right-padded language tokens, a
    padded range in the middle, stale rows behind the prefix keys in both caches, key ranges of 2, 3 and 4 blocks of 256,
    sms < items: the persistent loop over many items per CTA (barrier parities, TMEM reuse)."""
    g = torch.Generator().manual_seed(11)
    q = (torch.randn((B * T, 8 * 256), generator=g) * 0.8).bfloat16()
    kc = torch.randn((B, pitch, 256), generator=g).bfloat16()
    vt = torch.randn((B, 256, pitch), generator=g).bfloat16()
    pad = torch.ones((B, P), dtype=torch.uint8)
    pad[0, P - 17:] = 0
    if B > 1:
        pad[1, 5:300] = 0
    att = torch.full((B * T, 8 * 256), 7.0, dtype=torch.bfloat16, device=cuda)
    dq, dk, dv, dp = q.to(cuda), kc.to(cuda), vt.to(cuda), pad.to(cuda)
    for _ in range(2):
        pb.check(pb.lib().pi0_attention_long(pb.ptr(dq), pb.ptr(dk), pb.ptr(dv), pb.ptr(att), pb.ptr(dp), None, B, T, P, 0,
                                             pitch, pitch, sms, _stream()), "pi0_attention_long")
    torch.cuda.synchronize()
    qf = q.float().view(B, T * 8, 256)
    sc = (qf @ kc[:, :P].float().transpose(1, 2)).bfloat16().float()
    sc = (sc * 0.0625).bfloat16().float()
    sc = torch.where(pad[:, None, :].bool(), sc, torch.tensor(NEG))
    pr = torch.softmax(sc, dim=-1).bfloat16().float()
    ref = (pr @ vt[:, :, :P].float().transpose(1, 2)).bfloat16().float().view(B * T, 8 * 256)
    assert torch.isfinite(att.float()).all()
    err = (att.float().cpu() - ref).abs().max().item()
    assert err <= 2 ** -6 * max(1.0, ref.abs().max().item()), err


@pytest.mark.parametrize("B,S,P", [(4, 17, 584), (2, 5, 328), (1, 51, 840), (70, 17, 584)])
def test_suffix_mode(cuda, B, S, P):
    g = torch.Generator().manual_seed(13 + S)
    n = P + S
    kv_pitch, vt_pitch = n, (n + 7) // 8 * 8
    q = (torch.randn((B * S, 8 * 256), generator=g) * 0.8).bfloat16()
    kc = torch.randn((B, kv_pitch, 256), generator=g).bfloat16()
    vt = torch.randn((B, 256, vt_pitch), generator=g).bfloat16()
    pad = (torch.rand((B, P), generator=g) > 0.15).to(torch.uint8)
    pad[:, 0] = 1
    pad[0, P - 30:] = 0
    cum = torch.tensor([1] + [2] * (S - 1), dtype=torch.int32)
    att = torch.full((B * S, 8 * 256), 7.0, dtype=torch.bfloat16, device=cuda)
    dq, dk, dv, dp, dc = q.to(cuda), kc.to(cuda), vt.to(cuda), pad.to(cuda), cum.to(cuda)
    pb.check(pb.lib().pi0_attention_long(pb.ptr(dq), pb.ptr(dk), pb.ptr(dv), pb.ptr(att), pb.ptr(dp), pb.ptr(dc), B, S, P, S,
                                         kv_pitch, vt_pitch, 148, _stream()), "pi0_attention_long")
    torch.cuda.synchronize()
    qf = q.float().view(B, S, 8, 256)
    sc = torch.einsum("bshd,bjd->bshj", qf, kc[:, :n].float()).bfloat16().float()
    sc = (sc * 0.0625).bfloat16().float()
    mask = torch.ones((B, S, n), dtype=torch.bool)
    mask[:, :, :P] = pad[:, None, :].bool()
    mask[:, :, P:] = (cum[None, :] <= cum[:, None])[None]
    sc = torch.where(mask[:, :, None, :], sc, torch.tensor(NEG))
    pr = torch.softmax(sc, dim=-1).bfloat16().float()
    ref = torch.einsum("bshj,bjd->bshd", pr, vt[:, :, :n].float().transpose(1, 2)).bfloat16().float().reshape(B * S, 8 * 256)
    assert torch.isfinite(att.float()).all()
    err = (att.float().cpu() - ref).abs().max().item()
    assert err <= 2 ** -6 * max(1.0, ref.abs().max().item()), err
    if S <= 48 and n <= 640:   # the kernel it replaces at these shapes
        att2 = torch.zeros_like(att)
        pb.check(pb.lib().pi0_decode_attention(pb.ptr(dq), pb.ptr(dk), pb.ptr(dv), pb.ptr(att2), pb.ptr(dp), pb.ptr(dc), B, S, P,
                                               kv_pitch, vt_pitch, 148, _stream()), "pi0_decode_attention")
        torch.cuda.synchronize()
        assert (att.float() - att2.float()).abs().max().item() <= 2 ** -6 * max(1.0, ref.abs().max().item())
