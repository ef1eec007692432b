"""Batch-1 (or small batch) latency of the full-depth path in CUDA-graph replay + per-label breakdown of one eager pass.
Development aid; run on the GPU box:  [PI0_MEGA=1] python tools/lat_b1.py [batch] [workload]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from int_act_b200.engine import Pi0Engine  # noqa: E402
from int_act_b200.weights import Dims, PackedWeights, random_state_dict  # noqa: E402
from pi0_helpers import synthetic_inputs  # noqa: E402

WORK = {"bridge": (1, 72, 4), "rephrase": (1, 64, 4), "libero": (2, 72, 16)}


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    n_cams, lang_len, horizon = WORK[sys.argv[2] if len(sys.argv) > 2 else "bridge"]
    dev = torch.device("cuda:0")
    dims = Dims(vocab=4096)
    sd = random_state_dict(dims, device=dev, seed=0)
    eng = Pi0Engine(PackedWeights(sd, dev, torch.bfloat16), n_cams=n_cams, lang_len=lang_len, horizon=horizon, max_batch=B)
    del sd
    inp = synthetic_inputs(B, n_cams, lang_len, horizon, vocab=dims.vocab)
    args = ([i.to(dev) for i in inp[0]], [m.to(dev) for m in inp[1]], inp[2].to(dev), inp[3].to(dev), inp[4].to(dev),
            inp[5].to(dev))
    eng.sample_actions(*args)
    lat = []
    for i in range(60):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        eng._static(B)["x_t"].normal_()
        a.record()
        eng._run_static(B)
        b.record()
        torch.cuda.synchronize()
        if i >= 10:
            lat.append(a.elapsed_time(b))
    lat.sort()
    prof = eng.profile_step(B)
    groups = {}
    for k, v in prof.items():
        groups[k.split(".")[0]] = groups.get(k.split(".")[0], 0.0) + v[0]
    print(json.dumps({"mega": os.environ.get("PI0_MEGA", "default"), "batch": B, "p50_ms": lat[len(lat) // 2],
                      "p99_ms": lat[-1], "launches": eng.launches_per_infer.get(B),
                      "phase_ms_profiled": {k: round(v, 3) for k, v in groups.items()},
                      "breakdown_ms": {k: [round(v[0], 4), v[1]] for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}}))


if __name__ == "__main__":
    main()
