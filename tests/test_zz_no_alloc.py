"""The allocation-free variant of the two-launch denoise attention (`pi0_decode_attention_split_static`: score scratch in a
static __device__ array inside the library, csrc/attn_split.cu) against the existing entry with caller-owned scratch
(`pi0_decode_attention_split`), through the C ABI, two calls in a row. The two share every kernel, so the outputs must be
bit-identical. Collected last on purpose (file name)."""
import pytest
import torch

from int_act_b200 import binding as pb

pytestmark = pytest.mark.gpu


def _stream():
    return pb.current_stream()


def _case(cuda, B, S, P, seed):
    g = torch.Generator().manual_seed(seed)
    n = P + S
    kv_pitch, vt_pitch = n, (n + 7) // 8 * 8
    q = (torch.randn((B * S, 8 * 256), generator=g) * 0.7).bfloat16().to(cuda)
    k = torch.randn((B, kv_pitch, 256), generator=g).bfloat16().to(cuda)
    vt = torch.randn((B, 256, vt_pitch), generator=g).bfloat16().to(cuda)
    pad = (torch.rand((B, P), generator=g) > 0.2).to(torch.uint8)
    pad[:, 0] = 1
    cum = torch.tensor(([1] + [2] * (S - 1))[:S], dtype=torch.int32)
    return q, k, vt, pad.to(cuda), cum.to(cuda), kv_pitch, vt_pitch


def _run(fn_name, cuda, case, B, S, P, with_scratch):
    q, k, vt, pad, cum, kv_pitch, vt_pitch = case
    att = torch.full((B * S, 8 * 256), float("nan"), dtype=torch.bfloat16, device=cuda)
    fn = getattr(pb.lib(), fn_name)
    if with_scratch:
        scratch = torch.zeros((B * 8 * S, (P + S + 31) // 32 * 32), dtype=torch.bfloat16, device=cuda)
        rc = fn(pb.ptr(q), pb.ptr(k), pb.ptr(vt), pb.ptr(att), pb.ptr(scratch), pb.ptr(pad), pb.ptr(cum), B, S, P, kv_pitch,
                vt_pitch, _stream())
    else:
        rc = fn(pb.ptr(q), pb.ptr(k), pb.ptr(vt), pb.ptr(att), pb.ptr(pad), pb.ptr(cum), B, S, P, kv_pitch, vt_pitch, _stream())
    pb.check(rc, fn_name)
    torch.cuda.synchronize()
    return att


def test_static_scratch_entry_matches_caller_scratch_entry_twice_in_a_row(cuda):
    # two different problems back to back: the second call finds the first call's scores in the shared array
    for B, S, P, seed in [(8, 5, 328, 21), (3, 6, 72, 22)]:
        case = _case(cuda, B, S, P, seed)
        ref = _run("pi0_decode_attention_split", cuda, case, B, S, P, with_scratch=True)
        out = _run("pi0_decode_attention_split_static", cuda, case, B, S, P, with_scratch=False)
        assert torch.isfinite(out.float()).all()
        assert torch.equal(out, ref), (out.float() - ref.float()).abs().max().item()
        assert ref.float().abs().max().item() > 0.05


def test_static_scratch_entry_rejects_sizes_beyond_its_maximum(cuda):
    max_bytes = int(pb.lib().pi0_decode_attention_split_static_max_bytes())
    assert max_bytes == 4 << 20
    B, S, P = 200, 5, 328            # 200 * 8 * 5 * 352 * 2 = 5.6 MB > 4 MiB
    assert B * 8 * S * 352 * 2 > max_bytes
    case = _case(cuda, 1, S, P, 23)   # never dereferenced: the size check comes first
    q, k, vt, pad, cum, kv_pitch, vt_pitch = case
    att = torch.zeros((S, 8 * 256), dtype=torch.bfloat16, device=cuda)
    rc = pb.lib().pi0_decode_attention_split_static(pb.ptr(q), pb.ptr(k), pb.ptr(vt), pb.ptr(att), pb.ptr(pad), pb.ptr(cum), B,
                                                     S, P, kv_pitch, vt_pitch, _stream())
    assert rc == -2
    msg = pb.lib().pi0_last_error().decode()
    assert "built-in buffer" in msg and "4194304" in msg
