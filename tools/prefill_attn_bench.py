"""Micro-benchmark (and, with -DPI0_TIMELINE, the per-tile timeline of CTA 0) of the fused prefix attention kernel."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from int_act_b200 import binding as pb  # noqa: E402

dev = torch.device("cuda:0")
lib = pb.lib()


def run(B, T=328, P=328, pitch=336, iters=20):
    q = (torch.randn((B * T, 2048), device=dev) * 0.8).bfloat16()
    kc = torch.randn((B, pitch, 256), device=dev).bfloat16()
    vt = torch.randn((B, 256, pitch), device=dev).bfloat16()
    pad = torch.ones((B, P), dtype=torch.uint8, device=dev)
    att = torch.zeros((B * T, 2048), dtype=torch.bfloat16, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    fn = lambda: pb.check(lib.pi0_prefill_attention(pb.ptr(q), pb.ptr(kc), pb.ptr(vt), pb.ptr(att), pb.ptr(pad), B, T, P,
                                                    pitch, pitch, 148, st), "prefill_attention")
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    tiles = B * ((T * 8 + 127) // 128)
    fl = 4.0 * B * T * 8 * P * 256
    msg = f"B={B:3d}: {us:8.1f} us  {fl / us / 1e6:7.1f} TFLOP/s  tiles/CTA {tiles / 148:5.2f}"
    if hasattr(lib, "pi0_debug_prefill_timeline"):
        buf = (C.c_ulonglong * 16)()
        lib.pi0_debug_prefill_timeline(buf)
        t = list(buf)
        names = ["S ready", "max done", "exp+sum done", "P written", "O ready", "O stored"]
        msg += " | first tile of CTA 0: " + ", ".join(f"{n} +{(t[i + 1] - t[0]) / 1e3:.2f}" for i, n in enumerate(names))
    print(msg, flush=True)


for B in (1, 8, 64):
    run(B)
