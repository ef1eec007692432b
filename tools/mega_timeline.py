"""Phase timeline of the persistent denoise kernel (layer 1 of the last step), needs the library built with
PI0_EXTRA_NVCC_FLAGS=-DPI0_TIMELINE. Development aid; run on the GPU box."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from int_act_b200 import binding as pb  # noqa: E402
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
for _ in range(3):
    eng._run_static(B)
torch.cuda.synchronize()
buf = (C.c_ulonglong * 128)()
lib = eng.lib
lib.pi0_debug_mega_timeline.argtypes = [C.c_void_p]
assert lib.pi0_debug_mega_timeline(buf) == 0
names = ["L start", "bar1 done", "norm staged", "qkv done", "bar2 done", "scores done", "bar2b done", "pv done", "bar3 done",
         "att staged", "o done", "bar4 done", "norm staged", "gu done", "bar5 done", "act staged", "down done"]
order = [0, 1, 2, 3, 4, 15, 16, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14]
for ci, cta in enumerate((0, 7, 64, 140)):
    t = [buf[ci * 32 + i] for i in order]
    print(f"cta {cta:3d}: " + " | ".join(f"{names[i]} +{(t[i] - t[0]) / 1e3:5.2f}" for i in range(1, len(order))))
