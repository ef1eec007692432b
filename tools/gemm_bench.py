"""Micro-benchmark of pi0_gemm over the shapes of the Pi0 path (development aid; run on the GPU box)."""
import sys
import torch
sys.path.insert(0, ".")
from int_act_b200 import binding as pb

dev = torch.device("cuda:0")


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def run(name, M, N, K, bn, bias=False, res=False, act=0, epi=0, alpha=1.0, impl=0, splits=1):
    A = torch.randn((M, K), device=dev, dtype=torch.bfloat16)
    W = torch.randn((N, K), device=dev, dtype=torch.bfloat16) * 0.02
    n_out = N // 2 if epi == pb.EPI_GEGLU else N
    out = torch.zeros((M, n_out), device=dev, dtype=torch.bfloat16)
    b = torch.randn((N,), device=dev) if bias else None
    r = torch.randn((M, n_out), device=dev, dtype=torch.bfloat16) if res else None
    ms = timeit(lambda: pb.gemm(A, W, out, bias=b, res=r, act=act, bn=bn, epi_mode=epi, alpha=alpha, impl=impl, splits=splits))
    print(f"{name:34s} impl={impl} M={M:6d} N={N:6d} K={K:6d} bn={bn:3d} bias={int(bias)} res={int(res)} act={act}: {ms*1e3:9.1f} us "
          f"{2.0*M*N*K/ms/1e9:8.1f} TFLOP/s", flush=True)


which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which == "pair":
    # single-CTA (impl 2) against the CTA-pair kernel (impl 3) on the many-tile shapes of the batch-64 path
    for impl in (2, 3):
        run("prefill gateup geglu", 20992, 32768, 2048, 256, epi=pb.EPI_GEGLU, impl=impl)
        run("prefill down +res", 20992, 2048, 16384, 256, res=True, impl=impl)
        run("prefill down +res 128", 20992, 2048, 16384, 128, res=True, impl=impl)
        run("prefill o_proj +res", 20992, 2048, 2048, 256, res=True, impl=impl)
        run("prefill qkv (plain)", 20992, 2560, 2048, 256, impl=impl)
        run("siglip fc1 bias gelu", 16384, 4304, 1152, 256, bias=True, act=1, impl=impl)
        run("siglip fc1 bias gelu 128", 16384, 4304, 1152, 128, bias=True, act=1, impl=impl)
        run("siglip fc2 bias res", 16384, 1152, 4304, 256, bias=True, res=True, impl=impl)
        run("siglip fc2 bias res 128", 16384, 1152, 4304, 128, bias=True, res=True, impl=impl)
        run("siglip qkv bias", 16384, 3456, 1152, 256, bias=True, impl=impl)
        run("siglip qkv bias 128", 16384, 3456, 1152, 128, bias=True, impl=impl)
        run("siglip out_proj bias res", 16384, 1152, 1152, 256, bias=True, res=True, impl=impl)
        run("siglip out_proj bias res 128", 16384, 1152, 1152, 128, bias=True, res=True, impl=impl)
        run("square 8192 k2048", 8192, 8192, 2048, 256, impl=impl)
    run("prefill down +res auto split", 20992, 2048, 16384, 256, res=True, impl=2, splits=0)
if which in ("all", "big"):
    run("gateup geglu", 20992, 32768, 2048, 256, epi=pb.EPI_GEGLU)
    run("plain 256", 20992, 16384, 2048, 256)
    run("plain 128", 20992, 16384, 2048, 128)
    run("down +res", 20992, 2048, 16384, 256, res=True)
    run("o_proj +res 256", 20992, 2048, 2048, 256, res=True)
    run("o_proj nores 256", 20992, 2048, 2048, 256)
    run("o_proj +res 128", 20992, 2048, 2048, 128, res=True)
    run("fc1 bias gelu 128", 16384, 4304, 1152, 128, bias=True, act=1)
    run("fc1 bias 128", 16384, 4304, 1152, 128, bias=True)
    run("fc1 plain 128", 16384, 4304, 1152, 128)
    run("fc1 plain 256", 16384, 4304, 1152, 256)
    run("fc1 plain 64", 16384, 4304, 1152, 64)
    run("fc2 bias res 128", 16384, 1152, 4304, 128, bias=True, res=True)
    run("qkv bias 256", 16384, 3456, 1152, 256, bias=True)
    run("qkv plain 128", 16384, 3456, 1152, 128)
    run("out_proj bias res 128", 16384, 1152, 1152, 128, bias=True, res=True)
    run("square 8192 k2048 256", 8192, 8192, 2048, 256)
    run("square 8192 k512 256", 8192, 8192, 512, 256)
    run("square 8192 k128 256", 8192, 8192, 128, 256)
if which in ("all", "small"):
    for M in (5, 20, 320):
        run("exp gateup geglu", M, 8192, 1024, 256, epi=pb.EPI_GEGLU)
        run("exp down", M, 1024, 4096, 32, res=True)
        run("exp down", M, 1024, 4096, 64, res=True)
        run("exp o_proj", M, 1024, 2048, 32, res=True)
        run("exp qkv (plain)", M, 2560, 1024, 32)
        run("exp qkv (plain)", M, 2560, 1024, 256)
