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
    scratch = torch.zeros((B * 8 * S, (n + 31) // 32 * 32), device=dev).bfloat16()

    def fn2():
        l = i[0] % L
        i[0] += 1
        pb.check(pb.lib().pi0_decode_attention_split(pb.ptr(q), pb.ptr(k[l]), pb.ptr(vt[l]), pb.ptr(att), pb.ptr(scratch),
                                                     pb.ptr(pad), pb.ptr(cum), B, S, P, n, vtp, pb.current_stream()))

    def fn3():
        l = i[0] % L
        i[0] += 1
        pb.check(pb.lib().pi0_decode_attention_batch(pb.ptr(q), pb.ptr(k[l]), pb.ptr(vt[l]), pb.ptr(att), pb.ptr(pad),
                                                     pb.ptr(cum), B, S, P, n, vtp, pb.current_stream()))
    print(f"B={B}: batch (smem-staged) {graph_time(fn3):.2f} us")
    if hasattr(pb.lib(), "pi0_debug_battn_timeline"):
        import ctypes as C
        buf = (C.c_ulonglong * 16)()
        pb.lib().pi0_debug_battn_timeline.argtypes = [C.c_void_p]
        pb.lib().pi0_debug_battn_timeline(buf)
        names = ["entry", "pdl wait", "K/Q issued", "K/Q landed", "scores", "softmax", "V landed", "PV+store"]
        print("   timeline: " + " | ".join(f"{names[i]} +{(buf[i] - buf[0]) / 1e3:.2f}" for i in range(1, 8)) + f" || V issued +{(buf[8]-buf[0])/1e3:.2f} masks ready +{(buf[9]-buf[0])/1e3:.2f}")
    print(f"B={B}: fused {graph_time(fn):.2f} us, split {graph_time(fn2):.2f} us per attention (rotating over {L} layers of KV)")
