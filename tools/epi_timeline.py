"""Epilogue duration of the last tile CTA 0 drained (tfull seen -> epilogue done), single-CTA tcgen05 kernel; needs the
library built with PI0_EXTRA_NVCC_FLAGS=-DPI0_TIMELINE. Development aid."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from int_act_b200 import binding as pb  # noqa: E402

dev = torch.device("cuda:0")
lib = pb.lib()


def run(name, M, N, K, bn, bias=False, res=False, act=0, epi=0):
    A = torch.randn((M, K), device=dev, dtype=torch.bfloat16)
    W = torch.randn((N, K), device=dev, dtype=torch.bfloat16) * 0.02
    n_out = N // 2 if epi == pb.EPI_GEGLU else N
    out = torch.zeros((M, n_out), device=dev, dtype=torch.bfloat16)
    b = torch.randn((N,), device=dev) if bias else None
    r = torch.randn((M, n_out), device=dev, dtype=torch.bfloat16) if res else None
    for _ in range(3):
        pb.gemm(A, W, out, bias=b, res=r, act=act, bn=bn, epi_mode=epi, impl=2)
    torch.cuda.synchronize()
    buf = (C.c_ulonglong * 16)()
    assert lib.pi0_debug_timeline(buf) == 0
    t = list(buf)
    tiles = ((M + 127) // 128) * ((N + bn - 1) // bn)
    print(f"{name:28s} tiles/CTA {tiles/148:5.1f} kernel {(t[7]-t[0])/1e3:8.1f} us  per tile {(t[7]-t[0])/1e3/((tiles+147)//148):6.2f} us  "
          f"last tile: tfull->epilogue done {(t[6]-t[5])/1e3:6.2f} us", flush=True)


run("siglip fc1 bias gelu", 16384, 4304, 1152, 256, bias=True, act=1)
run("siglip fc1 plain", 16384, 4304, 1152, 256)
run("siglip qkv bias", 16384, 3456, 1152, 256, bias=True)
run("siglip out bias res", 16384, 1152, 1152, 256, bias=True, res=True)
run("siglip fc2 bias res", 16384, 1152, 4304, 256, bias=True, res=True)
run("gateup geglu", 20992, 32768, 2048, 256, epi=pb.EPI_GEGLU)
run("o_proj res", 20992, 2048, 2048, 256, res=True)
