"""fp32 regime GEMM on the tensor cores (csrc/gemm_x3.cu): error against an fp64 product and time, next to torch's fp32
matmul (cuBLAS SGEMM, TF32 off) and the SIMT check kernel (development aid)."""
import ctypes as C
import sys
import torch
sys.path.insert(0, ".")
from int_act_b200 import binding as pb
sys.path.insert(0, "tools")

dev = torch.device("cuda:0")
torch.backends.cuda.matmul.allow_tf32 = False


def timed(fn, n=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


for M, N, K in [(40, 1024, 4096), (40, 8192, 1024), (5, 2560, 1024), (328, 2048, 2048), (2624, 2048, 16384), (2624, 32768, 2048), (2048, 1152, 4304), (320, 1024, 4096),
                (20992, 2048, 16384)]:
    g = torch.Generator(device=dev).manual_seed(M + N + K)
    A = torch.randn((M, K), generator=g, device=dev)
    W = torch.randn((N, K), generator=g, device=dev) * K ** -0.5
    ref = (A[:256].double() @ W.double().t())
    out = torch.zeros((M, N), device=dev)
    d = pb.GemmDesc()
    d.A, d.W, d.out = pb.ptr(A), pb.ptr(W), pb.ptr(out)
    d.M, d.N, d.K, d.lda, d.ldw, d.ldc = M, N, K, K, K, N
    d.nz0 = d.nz1 = 1
    d.bn = 256
    d.alpha = 1.0
    d.vt_col0 = 0x7FFFFFFF
    d.vt_hd = d.vt_heads = d.vt_rpb = 1
    ba = int(pb.lib().pi0_gemm_f32x3_scratch_bytes(1, M, K))
    bw = int(pb.lib().pi0_gemm_f32x3_scratch_bytes(1, N, K))
    sa = torch.empty(ba, dtype=torch.uint8, device=dev)
    sw = torch.empty(bw, dtype=torch.uint8, device=dev)
    bp = 64 << 20
    sp = torch.empty(bp, dtype=torch.uint8, device=dev)

    def x3():
        pb.check(pb.lib().pi0_gemm_f32x3(C.byref(d), pb.ptr(sa), ba, pb.ptr(sw), bw, pb.ptr(sp), bp, 148, pb.current_stream()), "x3")
    t_x3 = timed(x3)
    e_x3 = (out[:256].double() - ref).abs().max().item()
    r_x3 = ((out[:256].double() - ref).abs().mean() / ref.abs().mean()).item()
    tt = timed(lambda: torch.matmul(A, W.t()))
    o2 = torch.matmul(A, W.t())
    e_t = (o2[:256].double() - ref).abs().max().item()
    r_t = ((o2[:256].double() - ref).abs().mean() / ref.abs().mean()).item()
    fl = 2.0 * M * N * K
    print(f"M={M} N={N} K={K}: x3 {t_x3*1e3:.0f} us ({fl/t_x3/1e9:.0f} TFLOP/s fp32-equivalent) max err {e_x3:.2e} mean rel {r_x3:.2e} | "
          f"torch fp32 {tt*1e3:.0f} us ({fl/tt/1e9:.0f} TFLOP/s) max err {e_t:.2e} mean rel {r_t:.2e}", flush=True)
