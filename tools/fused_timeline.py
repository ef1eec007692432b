"""Phase timeline of the fused denoise-step kernel (layer 1), needs the instrumented library (see tools/loop_timeline.py):
PI0_LIB=int-act_b200/libpi0b200_tl.so python tools/fused_timeline.py [batch]. Development aid."""
import ctypes as C
import os
import sys

import torch

os.environ.setdefault("PI0_FUSED", "1")   # the fused kernel is opt-in

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from int_act_b200.engine import Pi0Engine  # noqa: E402
from int_act_b200.weights import Dims, PackedWeights, random_state_dict  # noqa: E402

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dims = Dims(n_siglip_layers=1, vocab=4096)
sd = random_state_dict(dims, device=dev, seed=0)
eng = Pi0Engine(PackedWeights(sd, dev, torch.bfloat16), max_batch=B)
print(eng.path_report(B))
st = eng._static(B)
st["lang_masks"].fill_(1)
st["x_t"].normal_()
for _ in range(3):
    eng._run_static(B)
torch.cuda.synchronize()
buf = (C.c_ulonglong * 128)()
lib = eng.lib
lib.pi0_debug_fused_timeline.argtypes = [C.c_void_p]
assert lib.pi0_debug_fused_timeline(buf) == 0
names = ["start", "qkv gemm", "bar", "rope reduce", "bar", "attention", "bar", "o gemm", "bar", "res+norm", "bar", "gu gemm", "bar",
         "geglu", "bar", "down gemm", "bar", "res+norm", "bar"]
for ci, cta in enumerate((0, 7, 64, 140)):
    t = [buf[ci * 32 + i] for i in range(19)]
    print(f"cta {cta:3d}: " + " | ".join(f"{names[i]} {(t[i] - t[i - 1]) / 1e3:5.2f}" for i in range(1, 19)) +
          f" | total {(t[18] - t[0]) / 1e3:6.2f} us")
