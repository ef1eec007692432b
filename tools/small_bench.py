"""Graph-replay micro-benchmark of the skinny GEMMs (denoise / batch-1 shapes): N back-to-back launches of one
pi0_gemm are captured in a CUDA graph so the number is GPU time per launch including the inter-kernel gap, not the
host launch rate. Development aid; run on the GPU box:  python tools/small_bench.py [filter]"""
import sys

import torch

sys.path.insert(0, ".")
from int_act_b200 import binding as pb  # noqa: E402

dev = torch.device("cuda:0")
REP = 40


def graph_time(fn):
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(REP):
            fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 5 / REP * 1e3  # us per launch


def run(name, M, N, K, bn, splits, epi=0, res=True):
    A = torch.randn((M, K), device=dev, dtype=torch.bfloat16)
    W = torch.randn((N, K), device=dev, dtype=torch.bfloat16) * 0.02
    n_out = N // 2 if epi == pb.EPI_GEGLU else N
    out = torch.zeros((M, n_out), device=dev, dtype=torch.bfloat16)
    r = torch.randn((M, n_out), device=dev, dtype=torch.bfloat16) if (res and epi == 0) else None
    us = graph_time(lambda: pb.gemm(A, W, out, res=r, bn=bn, splits=splits, epi_mode=epi))
    wbytes = N * K * 2
    print(f"{name:22s} M={M:5d} N={N:6d} K={K:6d} bn={bn:3d} sp={splits}: {us:8.2f} us  "
          f"{wbytes / us / 1e3:7.1f} GB/s(W)  {2.0 * M * N * K / us / 1e6:8.1f} TFLOP/s", flush=True)


def main():
    flt = sys.argv[1] if len(sys.argv) > 1 else ""
    x = torch.zeros(64, device=dev)
    y = torch.zeros(64, device=dev, dtype=torch.bfloat16)
    print(f"floor (torch 64-elem copy kernel): {graph_time(lambda: y.copy_(x)):.2f} us/launch")

    print("max active clusters by size:", {sp: pb.lib().pi0_debug_max_clusters(sp) for sp in range(2, 9)})
    cases = []
    for M in (5, 320):
        for bn, sp in ((32, 1), (64, 1), (64, 4), (64, 8), (128, 4), (128, 6), (128, 8), (256, 4), (256, 8), (0, 0)):
            cases.append(("exp down", M, 1024, 4096, bn, sp, 0))
        for bn, sp in ((32, 1), (64, 8), (128, 6), (256, 8), (0, 0)):
            cases.append(("exp o_proj", M, 1024, 2048, bn, sp, 0))
        for bn, sp in ((256, 1), (256, 2), (256, 4), (256, 0)):
            cases.append(("exp qkv(plain)", M, 2560, 1024, bn, sp, 0))
        for sp in (1, 2, 4, 0):
            cases.append(("exp gateup geglu", M, 8192, 1024, 256, sp, pb.EPI_GEGLU))
    for bn, sp in ((64, 1), (256, 4), (256, 6), (128, 2), (0, 0)):
        cases.append(("b1 prefill down", 328, 2048, 16384, bn, sp, 0))
    for bn, sp in ((256, 1), (128, 1), (0, 0)):
        cases.append(("b1 prefill gateup", 328, 32768, 2048, 256, sp if bn == 256 else 0, pb.EPI_GEGLU))
    for bn, sp in ((64, 1), (256, 4), (0, 0)):
        cases.append(("b1 prefill o_proj", 328, 2048, 2048, bn, sp, 0))
    for bn, sp in ((128, 1), (256, 2), (256, 4), (0, 0)):
        cases.append(("b1 siglip fc1", 256, 4304, 1152, bn, sp, 0))
    for bn, sp in ((64, 1), (256, 8), (128, 6), (0, 0)):
        cases.append(("b1 siglip fc2", 256, 1152, 4304, bn, sp, 0))
    for c in cases:
        if flt in c[0]:
            run(*c[:6], epi=c[6])



if __name__ == "__main__":
    main()
