"""One eager (non-graph) `pi0_infer` pass at a given batch on the full bridge architecture, for ncu launch lists and
single-kernel captures (run on the GPU box: see profiles/README.md for the exact command lines)."""
import sys

import torch

sys.path.insert(0, ".")
from int_act_b200.engine import Pi0Engine
from int_act_b200.weights import Dims, PackedWeights, random_state_dict

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
passes = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda:0")
dims = Dims(vocab=8192)
sd = random_state_dict(dims, device=dev, seed=0)
eng = Pi0Engine(PackedWeights(sd, dev, torch.bfloat16, free_source=True), max_batch=B, use_cuda_graph=False)
del sd
st = eng._static(B)
g = torch.Generator(device=dev).manual_seed(1)
st["images"][0].uniform_(0, 1, generator=g)
st["tokens"].random_(3, dims.vocab, generator=g)
st["lang_masks"].fill_(1)
st["state"][:, :7].normal_(0, 0.5, generator=g)
for _ in range(passes):
    st["x_t"].normal_(generator=g)
    eng.lib.pi0_launch_count(1)
    eng._run_static(B)
    n_launch = int(eng.lib.pi0_launch_count(1))
torch.cuda.synchronize()
print("ok", float(st["x_t"].abs().max()))
print("launches_per_pass", n_launch)
