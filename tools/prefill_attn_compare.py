"""Prefix-pass attention at the bridge / rephrase shapes: the one-tile kernel (attn_prefill.cu) against the blocked two-pass
kernel (attn_long.cu), graph-replay time per launch (development aid)."""
import sys
import torch
sys.path.insert(0, ".")
from int_act_b200 import binding as pb
sys.path.insert(0, "tools")
from small_bench import graph_time  # noqa

dev = torch.device("cuda:0")
for B, P in [(64, 328), (64, 320), (32, 328), (8, 328)]:
    T = P
    pitch = (P + 5 + 7) // 8 * 8
    L = 4
    q = (torch.randn((B * T, 2048), device=dev) * 0.8).bfloat16()
    k = torch.randn((L, B, pitch, 256), device=dev).bfloat16()
    vt = torch.randn((L, B, 256, pitch), device=dev).bfloat16()
    att1 = torch.zeros_like(q)
    att2 = torch.zeros_like(q)
    pad = torch.ones((B, P), dtype=torch.uint8, device=dev)
    pad[:, P - 20:] = 0
    i = [0]

    def f1():
        i[0] += 1
        l = i[0] % L
        pb.check(pb.lib().pi0_prefill_attention(pb.ptr(q), pb.ptr(k[l]), pb.ptr(vt[l]), pb.ptr(att1), pb.ptr(pad), B, T, P, pitch,
                                                pitch, 148, pb.current_stream()), "prefill")

    def f2():
        i[0] += 1
        l = i[0] % L
        pb.check(pb.lib().pi0_attention_long(pb.ptr(q), pb.ptr(k[l]), pb.ptr(vt[l]), pb.ptr(att2), pb.ptr(pad), None, B, T, P, 0,
                                             pitch, pitch, 148, pb.current_stream()), "long")
    t1, t2 = graph_time(f1), graph_time(f2)
    i[0] = 0
    f1(); i[0] = 0; f2()
    torch.cuda.synchronize()
    print(f"B={B} P={P}: one-tile kernel {t1:.1f} us, blocked two-pass kernel {t2:.1f} us; max diff "
          f"{(att1.float() - att2.float()).abs().max().item():.3e}", flush=True)
