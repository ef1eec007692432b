"""Denoise-step attention at the bridge shape (5 suffix tokens over 333 keys): batch kernel (attn_batch.cu) against the blocked
two-pass tcgen05 kernel (attn_long.cu), graph-replay time per launch, rotating over 18 layers of KV (development aid)."""
import sys
import torch
sys.path.insert(0, ".")
from int_act_b200 import binding as pb
sys.path.insert(0, "tools")
from small_bench import graph_time  # noqa

dev = torch.device("cuda:0")
for B in (16, 64, 128, 256):
    S, P = 5, 328
    n = P + S
    vtp = (n + 7) // 8 * 8
    L = 18
    q = torch.randn((B * S, 2048), device=dev).bfloat16()
    k = torch.randn((L, B, n, 256), device=dev).bfloat16()
    vt = torch.randn((L, B, 256, vtp), device=dev).bfloat16()
    a1, a2 = torch.zeros_like(q), torch.zeros_like(q)
    pad = torch.ones((B, P), dtype=torch.uint8, device=dev)
    cum = torch.tensor([1, 2, 2, 2, 2], dtype=torch.int32, device=dev)
    i = [0]

    def f1():
        l = i[0] % L
        i[0] += 1
        pb.check(pb.lib().pi0_decode_attention_batch(pb.ptr(q), pb.ptr(k[l]), pb.ptr(vt[l]), pb.ptr(a1), pb.ptr(pad), pb.ptr(cum),
                                                     B, S, P, n, vtp, pb.current_stream()), "batch")

    def f2():
        l = i[0] % L
        i[0] += 1
        pb.check(pb.lib().pi0_attention_long(pb.ptr(q), pb.ptr(k[l]), pb.ptr(vt[l]), pb.ptr(a2), pb.ptr(pad), pb.ptr(cum), B, S, P,
                                             S, n, vtp, 148, pb.current_stream()), "long")
    t1, t2 = graph_time(f1), graph_time(f2)
    i[0] = 0; f1(); i[0] = 0; f2()
    torch.cuda.synchronize()
    print(f"B={B}: batch kernel {t1:.1f} us, blocked two-pass tcgen05 kernel {t2:.1f} us; max diff "
          f"{(a1.float() - a2.float()).abs().max().item():.3e}", flush=True)
