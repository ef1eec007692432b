"""In-kernel phase timeline of one tcgen05 GEMM launch inside a graph-replayed chain (needs the library built with
PI0_EXTRA_NVCC_FLAGS=-DPI0_TIMELINE). Development aid."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from int_act_b200 import binding as pb  # noqa: E402

dev = torch.device("cuda:0")
lib = pb.lib()
NAMES = ["entry", "setup done", "first TMA issued", "first full", "last commit", "tfull seen", "epilogue done", "exit"]


def run(M, N, K, bn, splits=1, rep=20):
    A = torch.randn((M, K), device=dev, dtype=torch.bfloat16)
    W = torch.randn((N, K), device=dev, dtype=torch.bfloat16) * 0.02
    out = torch.zeros((M, N), device=dev, dtype=torch.bfloat16)
    fn = lambda: pb.gemm(A, W, out, bn=bn, splits=splits)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(rep):
            fn()
    g.replay(); g.replay()
    torch.cuda.synchronize()
    buf = (C.c_ulonglong * 16)()
    assert lib.pi0_debug_timeline(buf) == 0
    t = list(buf)
    print(f"M={M} N={N} K={K} bn={bn}: prev-exit->entry {(t[0]-t[8])/1e3:6.2f} us | " +
          " | ".join(f"{NAMES[i]} +{(t[i]-t[0])/1e3:5.2f}" for i in range(1, 8)) +
          f" | kb8 +{(t[9]-t[0])/1e3:5.2f} kb16 +{(t[10]-t[0])/1e3:5.2f} kb32 +{(t[11]-t[0])/1e3:5.2f} last-stage-first-fill +{(t[12]-t[0])/1e3:5.2f} (issued +{(t[13]-t[0])/1e3:5.2f}) producer cycles wait/expect/tmaA/tmaB = {t[14]>>48}/{(t[14]>>32)&0xffff}/{(t[14]>>16)&0xffff}/{t[14]&0xffff}; mma-thread wait/fence/4xmma/commit = {t[15]>>48}/{(t[15]>>32)&0xffff}/{(t[15]>>16)&0xffff}/{t[15]&0xffff}")


run(5, 1024, 4096, 64)
run(128, 1024, 4096, 64)
run(5, 1024, 4096, 256)
run(5, 1024, 64, 64)
run(5, 8192, 1024, 256)
run(320, 1024, 4096, 32)
run(5, 256, 64, 256)
