"""Per-label time of one fp32-regime infer (eng.profile_step) at batch B (development aid)."""
import sys
import torch
sys.path.insert(0, ".")
from int_act_b200.engine import Pi0Engine
from int_act_b200.weights import Dims, PackedWeights, random_state_dict

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = torch.device("cuda:0")
dims = Dims(vocab=8192)
sd = random_state_dict(dims, device=dev, seed=0)
eng = Pi0Engine(PackedWeights(sd, dev, torch.float32, free_source=True), max_batch=B, dtype="fp32", use_cuda_graph=False)
del sd
st = eng._static(B)
st["images"][0].uniform_(0, 1)
st["tokens"].random_(3, dims.vocab)
st["lang_masks"].fill_(1)
st["x_t"].normal_()
eng._run_static(B)
torch.cuda.synchronize()
prof = eng.profile_step(B)
tot = sum(v[0] for v in prof.values())
print(f"batch {B}: {tot:.1f} ms summed")
for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    print(f"  {k:28s} {v[0]:8.3f} ms  x{v[1]}")
