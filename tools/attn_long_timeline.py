"""In-kernel timeline of the long-key attention (second item of CTA 0, first softmax thread) + time per launch in graph replay.
Build: PI0_EXTRA_NVCC_FLAGS=-DPI0_TIMELINE python int-act_b200/build.py --variant=tl ; run with PI0_LIB=int-act_b200/libpi0b200_tl.so"""
import ctypes as C
import sys
import torch
sys.path.insert(0, ".")
from int_act_b200 import binding as pb
sys.path.insert(0, "tools")
from small_bench import graph_time  # noqa

dev = torch.device("cuda:0")
for mode, B, T, P, S in [("prefix", 64, 584, 584, 0), ("suffix", 64, 17, 584, 17)]:
    n = P + S
    kvp, vtp = n, (n + 7) // 8 * 8
    q = (torch.randn((B * T, 2048), device=dev) * 0.8).bfloat16()
    k = torch.randn((B, kvp, 256), device=dev).bfloat16()
    vt = torch.randn((B, 256, vtp), device=dev).bfloat16()
    att = torch.zeros_like(q)
    pad = torch.ones((B, P), dtype=torch.uint8, device=dev)
    cum = torch.tensor([1] + [2] * max(S - 1, 0), dtype=torch.int32, device=dev)

    def fn():
        pb.check(pb.lib().pi0_attention_long(pb.ptr(q), pb.ptr(k), pb.ptr(vt), pb.ptr(att), pb.ptr(pad), pb.ptr(cum) if S else None,
                                             B, T, P, S, kvp, vtp, 148, pb.current_stream()), "attention_long")
    t = graph_time(fn)
    items = B * ((T * 8 + 127) // 128)
    print(f"{mode}: {t:.1f} us per launch, {items} items, {t / max(1, -(-items // 148)):.1f} us per item wave")
    if hasattr(pb.lib(), "pi0_debug_attn_long_timeline"):
        buf = (C.c_ulonglong * 64)()
        pb.lib().pi0_debug_attn_long_timeline.argtypes = [C.c_void_p]
        pb.lib().pi0_debug_attn_long_timeline(buf)
        t0 = buf[0]
        names = {0: "item start", 20: "exchange done", 40: "O ready", 41: "item done"}
        for kb in range(3):
            names.update({1 + kb * 4: f"p1 b{kb} mask", 2 + kb * 4: f"p1 b{kb} S ready", 3 + kb * 4: f"p1 b{kb} scores",
                          4 + kb * 4: f"p1 b{kb} exp", 21 + kb * 4: f"p2 b{kb} scores", 22 + kb * 4: f"p2 b{kb} P free",
                          23 + kb * 4: f"p2 b{kb} P written"})
        prev = t0
        for i in sorted(names):
            if buf[i]:
                print(f"   {names[i]:18s} +{(buf[i] - t0) / 1e3:7.2f} us (step {(buf[i] - prev) / 1e3:6.2f})")
                prev = buf[i]
