"""Per-phase CUDA-graph timing of the Pi0 path (development aid; run on the GPU box):
    python tools/phase_bench.py [batch ...]
Each phase (SigLIP, language embed + prefill, denoise loop) is captured in its own CUDA graph over the engine's static
buffers, so the numbers include programmatic-dependent-launch overlap exactly as the product graph does (the
per-launch event profile of bench.py serialises launches and overstates the small kernels)."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from int_act_b200 import binding as pb  # noqa: E402
from int_act_b200.engine import Pi0Engine  # noqa: E402
from int_act_b200.weights import Dims, PackedWeights, random_state_dict  # noqa: E402

dev = torch.device("cuda:0")


def graph_ms(fn, reps=10):
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    batches = [int(a) for a in sys.argv[1:]] or [1, 64]
    dims = Dims()
    sd = random_state_dict(dims, device=dev, seed=0)
    packed = PackedWeights(sd, dev, torch.bfloat16)
    del sd
    torch.cuda.empty_cache()
    eng = Pi0Engine(packed, max_batch=max(batches))
    lib = eng.lib
    for B in batches:
        st = eng._static(B)
        st["images"][0].uniform_(0, 1)
        st["tokens"].random_(3, 257152)
        st["lang_masks"].fill_(1)
        st["x_t"].normal_()
        s = pb.current_stream

        def siglip():
            pb.check(lib.pi0_siglip_forward(eng.ctx, B, 0, pb.ptr(st["images"][0]), 2.0, -1.0, s()), "siglip")

        def prefill():
            pb.check(lib.pi0_embed_language(eng.ctx, B, pb.ptr(st["tokens"]), s()), "embed")
            pb.check(lib.pi0_prefill(eng.ctx, B, pb.ptr(st["img_masks"]), pb.ptr(st["lang_masks"]), s()), "prefill")

        def denoise():
            pb.check(lib.pi0_denoise_loop(eng.ctx, B, pb.ptr(st["state"]), pb.ptr(st["x_t"]), s()), "denoise")

        def whole():
            siglip()
            prefill()
            denoise()

        r = {}
        for name, fn in (("siglip", siglip), ("prefill", prefill), ("denoise", denoise), ("whole", whole)):
            st["x_t"].normal_()
            r[name] = graph_ms(fn)
        print(f"B={B:3d}  siglip {r['siglip']:8.3f} ms  prefill {r['prefill']:8.3f} ms  denoise {r['denoise']:8.3f} ms  "
              f"sum {r['siglip'] + r['prefill'] + r['denoise']:8.3f}  whole {r['whole']:8.3f} ms", flush=True)


if __name__ == "__main__":
    main()
