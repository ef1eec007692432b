"""Per-tile timeline of CTA pair 0 of the cta_group::2 GEMM (MMA first issue / last commit in the leader, tfull seen /
drained in both CTAs' epilogues); needs PI0_EXTRA_NVCC_FLAGS=-DPI0_TIMELINE. Development aid."""
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
        pb.gemm(A, W, out, bias=b, res=r, act=act, bn=bn, epi_mode=epi, impl=3)
    torch.cuda.synchronize()
    buf = (C.c_ulonglong * 128)()
    assert lib.pi0_debug_pair_timeline(buf) == 0
    t = list(buf)
    t0 = t[0]
    print(name)
    for tile in range(8):
        l = [t[(0 * 16 + tile) * 4 + k] for k in range(4)]
        p = [t[(1 * 16 + tile) * 4 + k] for k in range(4)]
        f = lambda v: f"{(v - t0) / 1e3:7.2f}"
        print(f"  tile {tile}: mma first {f(l[0])} last {f(l[1])} | leader epi {f(l[2])} -> {f(l[3])} | peer epi {f(p[2])} -> {f(p[3])}")


run("siglip fc1 bias gelu", 16384, 4304, 1152, 256, bias=True, act=1)
run("siglip qkv bias", 16384, 3456, 1152, 256, bias=True)
run("gateup geglu", 20992, 32768, 2048, 256, epi=pb.EPI_GEGLU)
