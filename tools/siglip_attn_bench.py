"""SigLIP attention: tcgen05 kernel (attn_siglip_tc.cu) vs the mma.sync kernel (attn_siglip.cu) - agreement with a torch
restatement and graph-replay time per launch, rotating over 4 q|k|v buffers (development aid)."""
import os
import sys
import torch
sys.path.insert(0, ".")
from int_act_b200 import binding as pb
sys.path.insert(0, "tools")
from small_bench import graph_time  # noqa

dev = torch.device("cuda:0")
T, H, D = 256, 16, 72


def ref_attention(qkv, B):
    q = qkv[:, : H * D].float().view(B, T, H, D).transpose(1, 2)
    k = qkv[:, H * D: 2 * H * D].float().view(B, T, H, D).transpose(1, 2)
    v = qkv[:, 2 * H * D:].float().view(B, T, H, D).transpose(1, 2)
    sc = (q @ k.transpose(2, 3)).bfloat16().float()
    sc = (sc * torch.tensor(72 ** -0.5, dtype=torch.float32, device=qkv.device)).bfloat16().float()
    pr = torch.softmax(sc, dim=-1).bfloat16().float()
    return (pr @ v).bfloat16().float().transpose(1, 2).reshape(B * T, H * D)


for B in (1, 3, 12, 64):
    g = torch.Generator(device=dev).manual_seed(3 + B)
    bufs = [(torch.randn((B * T, 3 * H * D), generator=g, device=dev) * 1.5).bfloat16() for _ in range(4)]
    att = torch.zeros((B * T, H * D), dtype=torch.bfloat16, device=dev)
    out = {}
    for mode in ("mma", "tc"):
        os.environ["PI0_SIGLIP_ATTN"] = mode
        att.zero_()
        pb.check(pb.lib().pi0_siglip_attention(pb.ptr(bufs[0]), pb.ptr(att), B, pb.current_stream()), "siglip_attention")
        torch.cuda.synchronize()
        out[mode] = att.clone()
        i = [0]

        def fn():
            i[0] += 1
            pb.check(pb.lib().pi0_siglip_attention(pb.ptr(bufs[i[0] % 4]), pb.ptr(att), B, pb.current_stream()), "x")
        t = graph_time(fn)
        ref = ref_attention(bufs[0], B)
        err = (out[mode].float() - ref).abs().max().item()
        flops = 2 * 2 * B * H * T * T * D
        print(f"B={B} {mode}: {t:.1f} us per launch ({flops / t / 1e6:.0f} TFLOP/s), max|err| vs torch {err:.3e} "
              f"(ref max {ref.abs().max().item():.2f})", flush=True)
    print(f"B={B}: tc vs mma max diff {(out['tc'].float() - out['mma'].float()).abs().max().item():.3e}", flush=True)
