"""SigLIP GEMM shapes on the CTA-pair kernel: 8 vs 12 epilogue warps (PI0_PAIR_EW3_MAX_K is read once per process, so run
twice:  PI0_PAIR_EW3_MAX_K=0 python tools/pair_ew_bench.py ; PI0_PAIR_EW3_MAX_K=2047 python tools/pair_ew_bench.py)."""
import os
import sys
sys.path.insert(0, ".")
sys.argv = [sys.argv[0], "none"]
exec(open("tools/gemm_bench.py").read().split("which = sys.argv[1]")[0])
print("PI0_PAIR_EW3_MAX_K =", os.environ.get("PI0_PAIR_EW3_MAX_K"))
for _ in range(2):
    run("siglip fc1 bias gelu", 16384, 4304, 1152, 256, bias=True, act=1, impl=3)
    run("siglip qkv bias", 16384, 3456, 1152, 256, bias=True, impl=3)
    run("siglip out_proj bias res", 16384, 1152, 1152, 256, bias=True, res=True, impl=3)
    run("prefill o_proj +res (K=2048)", 20992, 2048, 2048, 256, res=True, impl=3)
