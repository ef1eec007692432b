"""Graph-replay timing of the fused decode attention kernel (development aid)."""
import sys
import torch
sys.path.insert(0, ".")
from int_act_b200 import binding as pb
sys.path.insert(0, "tools")
from small_bench import graph_time  # noqa

dev = torch.device("cuda:0")
for B in (1, 8, 64):
    S, P = 5, 328
    n = P + S
    vtp = (n + 7) // 8 * 8
    L = 18
    q = torch.randn((B * S, 2048), device=dev).bfloat16()
    k = torch.randn((L, B, n, 256), device=dev).bfloat16()
    vt = torch.randn((L, B, 256, vtp), device=dev).bfloat16()
    att = torch.zeros_like(q)
    pad = torch.ones((B, P), dtype=torch.uint8, device=dev)
    cum = torch.tensor([1, 2, 2, 2, 2], dtype=torch.int32, device=dev)
    i = [0]

    def fn():
        l = i[0] % L
        i[0] += 1
        pb.check(pb.lib().pi0_decode_attention(pb.ptr(q), pb.ptr(k[l]), pb.ptr(vt[l]), pb.ptr(att), pb.ptr(pad),
                                               pb.ptr(cum), B, S, P, n, vtp, 148, pb.current_stream()))
    print(f"B={B}: {graph_time(fn):.2f} us per launch (rotating over {L} layers of KV)")
