"""Phase timeline of the whole-loop persistent denoise kernel (global layer 1 = layer 1 of step 0), needs the instrumented
library: PI0_EXTRA_NVCC_FLAGS=-DPI0_TIMELINE python int-act_b200/build.py --variant=tl, then on the GPU box
PI0_LIB=int-act_b200/libpi0b200_tl.so python tools/loop_timeline.py [batch]. Development aid."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from int_act_b200.engine import Pi0Engine  # noqa: E402
from int_act_b200.weights import Dims, PackedWeights, random_state_dict  # noqa: E402

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
dims = Dims(n_siglip_layers=1, vocab=4096)
sd = random_state_dict(dims, device=dev, seed=0)
eng = Pi0Engine(PackedWeights(sd, dev, torch.bfloat16), max_batch=B)
st = eng._static(B)
st["lang_masks"].fill_(1)
st["x_t"].normal_()
lib = eng.lib
lib.pi0_debug_loop_stats.argtypes = [C.c_void_p, C.c_int]
stats = (C.c_ulonglong * 4)()
for _ in range(3):
    eng._run_static(B)
torch.cuda.synchronize()
lib.pi0_debug_loop_stats(stats, 1)
eng._run_static(B)
torch.cuda.synchronize()
lib.pi0_debug_loop_stats(stats, 0)
print(f"one infer: gathers {stats[3]}, words that missed the first poll {stats[1]}, retry rounds {stats[2]}")
buf = (C.c_ulonglong * 128)()
lib.pi0_debug_loop_timeline.argtypes = [C.c_void_p]
assert lib.pi0_debug_loop_timeline(buf) == 0
names = {0: "L start", 11: "hintX", 1: "norm staged", 2: "qkv done", 12: "hintQ", 3: "scores done", 13: "hintSC", 4: "pv done",
         14: "hintATT", 5: "att staged", 6: "o done", 15: "hintY", 7: "norm staged", 8: "gu done", 16: "hintACT",
         9: "act staged", 10: "down done"}
order = [0, 11, 1, 2, 12, 3, 13, 4, 14, 5, 6, 15, 7, 8, 16, 9, 10]
for ci, cta in enumerate((0, 7, 64, 140)):
    t0 = buf[ci * 32]
    cells = []
    for i in order[1:]:
        v = buf[ci * 32 + i]
        cells.append(f"{names[i]} +{(v - t0) / 1e3:6.2f}" if v >= t0 and v else f"{names[i]}   --  ")
    print(f"cta {cta:3d}: " + " | ".join(cells))
