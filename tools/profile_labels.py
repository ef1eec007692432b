"""Per-label CUDA-event breakdown of one eager pass (development aid): python tools/profile_labels.py B [filter]"""
import sys

import torch

sys.path.insert(0, ".")
from int_act_b200.engine import Pi0Engine  # noqa: E402
from int_act_b200.weights import Dims, PackedWeights, random_state_dict  # noqa: E402

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
flt = sys.argv[2] if len(sys.argv) > 2 else ""
sd = random_state_dict(Dims(), device=dev, seed=0)
packed = PackedWeights(sd, dev, torch.bfloat16)
del sd
eng = Pi0Engine(packed, max_batch=B)
st = eng._static(B)
st["lang_masks"].fill_(1)
st["x_t"].normal_()
eng.profile_step(B)
prof = eng.profile_step(B)
for k, (ms, n) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    if flt in k:
        print(f"{k:28s} {ms:8.3f} ms  {n:5d} launches  {1e3 * ms / n:8.2f} us/launch")
