"""In-kernel timeline of the tcgen05 SigLIP attention (third item of CTA 0, first softmax thread). Build the instrumented
library first: PI0_EXTRA_NVCC_FLAGS=-DPI0_TIMELINE python int-act_b200/build.py --variant=tl ; run with
PI0_LIB=int-act_b200/libpi0b200_tl.so."""
import ctypes as C
import sys
import torch
sys.path.insert(0, ".")
from int_act_b200 import binding as pb

dev = torch.device("cuda:0")
B, T, H, D = 64, 256, 16, 72
qkv = (torch.randn((B * T, 3 * H * D), device=dev) * 1.5).bfloat16()
att = torch.zeros((B * T, H * D), dtype=torch.bfloat16, device=dev)
for _ in range(3):
    pb.check(pb.lib().pi0_siglip_attention(pb.ptr(qkv), pb.ptr(att), B, pb.current_stream()), "siglip_attention")
torch.cuda.synchronize()
buf = (C.c_ulonglong * 32)()
pb.lib().pi0_debug_siglip_tc_timeline.argtypes = [C.c_void_p]
pb.lib().pi0_debug_siglip_tc_timeline(buf)
names = ["h0 enter", "h0 S ready", "h0 scores", "h0 exp", "h0 exchange", "h0 P written",
         "h1 enter", "h1 S ready", "h1 scores", "h1 exp", "h1 exchange(+P free)", "h1 P written",
         "ep0 O ready", "ep0 done", "ep1 O ready", "ep1 done"]
t0 = buf[0]
prev = t0
for i, n in enumerate(names):
    print(f"{n:24s} +{(buf[i] - t0) / 1e3:7.2f} us  (step {(buf[i] - prev) / 1e3:6.2f})")
    prev = buf[i]
