"""One eager fp32-regime infer at batch B on the full architecture, for ncu launch lists (development aid)."""
import sys
import torch
sys.path.insert(0, ".")
from int_act_b200.engine import Pi0Engine
from int_act_b200.weights import Dims, PackedWeights, random_state_dict

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda:0")
dims = Dims(vocab=8192)
sd = random_state_dict(dims, device=dev, seed=0)
eng = Pi0Engine(PackedWeights(sd, dev, torch.float32, free_source=True), max_batch=B, dtype="fp32", use_cuda_graph=False)
del sd
st = eng._static(B)
st["images"][0].uniform_(0, 1)
st["tokens"].random_(3, dims.vocab)
st["lang_masks"].fill_(1)
for _ in range(passes):
    st["x_t"].normal_()
    eng.lib.pi0_launch_count(1)
    eng._run_static(B)
    n = int(eng.lib.pi0_launch_count(1))
torch.cuda.synchronize()
print("launches_per_pass", n)
