#!/usr/bin/env python
"""bench.py - Pi0 `infer_action` throughput / latency on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch 64] [--workload bridge]
  torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...      (one rank per GPU, weak scaling)

A "step" = one `infer_action` over one rollout batch of `--batch` synthetic Bridge-shaped observations PER GPU
(SigLIP -> Gemma-2B prefill into the KV cache -> 10-step flow-matching Euler loop), i.e. BASELINE config 3.
`value`   = action chunks/s, whole job, inputs already resident in HBM (CUDA-graph replay of the whole path).
`e2e`     = the same through PI0Policy.infer_action with pinned HOST buffers: H2D of the observation, the path, D2H of
            the chunk, every step.
`roofline`= the dominant kernel (prefill gate/up GeGLU tcgen05 GEMM) timed with CUDA events inside one step.
`latency_b1` = BASELINE's second number: p50/p99 batch-1 latency (config 2) and its HBM-roofline fraction.
`cpu_baseline` (N=1) and `--impl reference` time the CPU oracle (a port of the reference's PyTorch path; the reference's
own implementation lives in the un-vendored lerobot submodule and cannot run here) on the host cores.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# stdout carries exactly ONE JSON line. Libraries write there too (NCCL prints its "NCCL version ..." banner on stdout
# whatever NCCL_DEBUG_FILE says), so file descriptor 1 is pointed at stderr for the whole process and the JSON line is
# written to the saved descriptor of the real stdout at the end.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line: dict) -> None:
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


import torch  # noqa: E402

WORKLOADS = {
    # name: (n_cams, lang_len, horizon)
    "bridge": (1, 72, 4),      # BASELINE configs 2/3 (P = 328, S = 5)
    "rephrase": (1, 64, 4),    # config 4 (P = 320)
    "libero": (2, 72, 16),     # config 5 (P = 584, S = 17)
}


def flops_per_chunk(n_cams, lang_len, horizon, num_steps=10):
    """F = ncam*F_sig + F_gem(P) + num_steps*F_exp(S, P)."""
    P, S = n_cams * 256 + lang_len, 1 + horizon
    f_sig = 2 * 256 * (27 * 15_224_832 + 677_376 + 2_359_296) + 27 * 4 * 256 ** 2 * 1152
    f_gem = 2 * 1_981_808_640 * P + 147_456 * P * P
    f_exp = 2 * 311_427_072 * S + 147_456 * S * (P + S)
    return n_cams * f_sig + f_gem + num_steps * f_exp


def bytes_per_chunk_b1(n_cams, lang_len, horizon, num_steps=10):
    """weights once (SigLIP + Gemma) + per step (expert weights + KV read) + KV write, bf16."""
    P = n_cams * 256 + lang_len
    kv = 2 * 18 * P * 256 * 2
    return 2 * (414.4e6 + 1981.8e6) + num_steps * (2 * 311.4e6 + kv) + kv


def roofline_ms_b1(n_cams, lang_len, horizon, peaks, num_steps=10):
    """north_star's batch-1 bound, per phase: the slower of FLOPs at the tensor peak and weight + KV bytes at HBM
    bandwidth (the prefill of a 328-token prefix is 328 FLOP per weight byte: above the ~215 FLOP/B ridge, tensor-bound;
    SigLIP at 256 tokens sits on the ridge; the 5-row denoise steps are pure weight streaming)."""
    P, S = n_cams * 256 + lang_len, 1 + horizon
    kv = 2 * 18 * P * 256 * 2
    f_sig = n_cams * (2 * 256 * (27 * 15_224_832 + 677_376 + 2_359_296) + 27 * 4 * 256 ** 2 * 1152)
    f_gem = 2 * 1_981_808_640 * P + 147_456 * P * P
    f_exp = num_steps * (2 * 311_427_072 * S + 147_456 * S * (P + S))
    b_sig, b_gem, b_exp = 2 * 414.4e6, 2 * 1981.8e6 + kv, num_steps * (2 * 311.4e6 + kv)
    tf, bw = peaks["bf16_tflops"] * 1e12, peaks["hbm_gbs"] * 1e9   # burst figures: a 9 ms pass is not power-capped
    phases = {"siglip": (f_sig, b_sig), "prefill": (f_gem, b_gem), "denoise": (f_exp, b_exp)}
    out = {k: {"tensor_ms": 1e3 * f / tf, "hbm_ms": 1e3 * b / bw} for k, (f, b) in phases.items()}
    total = sum(max(v["tensor_ms"], v["hbm_ms"]) for v in out.values())
    return total, out


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.02)

    def result(self):
        self.stop_flag = True
        if self.is_alive():
            self.join(timeout=1.0)
        med = statistics.median(self.samples) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def synthetic_observations(B, n_cams, lang_len, seed, pinned):
    """Seeded Bridge-shaped observations at the policy boundary: images f32 [B,3,224,224] in
    [0,1], state [B,7] in [-1,1], token ids with bos=2 / trailing "\\n"=108, all text positions valid."""
    g = torch.Generator().manual_seed(seed)
    batch = {}
    keys = ["observation.images.top", "observation.images.wrist", "observation.images.cam2", "observation.images.cam3"]
    for c in range(n_cams):
        batch[keys[c]] = torch.rand((B, 3, 224, 224), generator=g)
    batch["observation.state"] = (torch.randn((B, 7), generator=g) * 0.5).clamp(-1, 1)
    tok = torch.randint(3, 257152, (B, lang_len), generator=g)
    tok[:, 0] = 2
    tok[:, -1] = 108
    batch["lang_tokens"] = tok
    batch["lang_masks"] = torch.ones((B, lang_len), dtype=torch.bool)
    if pinned:
        batch = {k: v.pin_memory() for k, v in batch.items()}
    return batch


def make_config(workload):
    from int_act_b200.config import bridge_config, libero_config

    n_cams, lang_len, horizon = WORKLOADS[workload]
    if workload == "libero":
        return libero_config()
    return bridge_config(tokenizer_max_length=lang_len)


# -----------------------------------------------------------------------------------------------------------------
# CPU oracle timing (cpu_baseline leg and the --impl reference arm)
# -----------------------------------------------------------------------------------------------------------------
def time_cpu_oracle(workload, n_chunks, warmup, log=None):
    """Times the fp32 CPU oracle (the shipped eval regime of the reference is fp32) on ONE observation
    of the workload per step. Returns (seconds per chunk list, cores)."""
    from int_act_b200.weights import Dims, random_state_dict
    from oracle.pi0_oracle import OracleDims, Pi0Oracle

    n_cams, lang_len, horizon = WORKLOADS[workload]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    dims = Dims(vocab=8192)  # the 1 GB embedding table is only gathered from (72 rows): a small table times the same
    gen_dev = "cuda" if torch.cuda.is_available() else "cpu"
    sd = random_state_dict(dims, device=gen_dev, seed=0)
    sd = {k: v.cpu() for k, v in sd.items()}
    od = OracleDims(n_cams=n_cams, lang_len=lang_len, horizon=horizon)
    orc = Pi0Oracle(od, sd, mode="fp32")
    del sd
    obs = synthetic_observations(1, n_cams, lang_len, seed=1234, pinned=False)
    images = [obs[k] for k in obs if k.startswith("observation.images")]
    masks = [torch.ones(1, dtype=torch.bool) for _ in images]
    tok = obs["lang_tokens"] % dims.vocab
    noise = torch.normal(0.0, 1.0, (1, horizon, 32), generator=torch.Generator().manual_seed(42))
    times = []
    for i in range(warmup + n_chunks):
        t0 = time.perf_counter()
        orc.sample_actions(images, masks, tok, obs["lang_masks"], obs["observation.state"], noise)
        dt = time.perf_counter() - t0
        if log:
            log(f"cpu oracle chunk {i}: {dt:.2f} s")
        if i >= warmup:
            times.append(dt)
    return times, cores


def time_gpu_eager(workload, B, dev, steps=3, b1_iters=20, log=None):
    """The reference-equivalent GPU number (BASELINE.md section 2 "GPU-eager"): the same op sequence the
    reference executes - PyTorch eager, torch.autocast(bfloat16), eager attention, one cuBLAS call per projection, the
    Python Euler loop - on this B200, same random-init weights / observation shapes / noise as the B200 arm
    (oracle/pi0_eager.py; the lerobot module itself is not in the tree). Device-timed with CUDA events."""
    from int_act_b200.weights import Dims, random_state_dict
    from oracle.pi0_eager import Pi0Eager
    from oracle.pi0_oracle import OracleDims

    n_cams, lang_len, horizon = WORKLOADS[workload]
    dims = Dims()
    sd = random_state_dict(dims, device=dev, seed=0)
    model = Pi0Eager(OracleDims(n_cams=n_cams, lang_len=lang_len, horizon=horizon), sd, dev, use_bf16=True)
    del sd
    torch.cuda.empty_cache()

    def inputs(b, seed):
        obs = synthetic_observations(b, n_cams, lang_len, seed=seed, pinned=False)
        imgs = [obs[k].to(dev) for k in obs if k.startswith("observation.images")]
        masks = [torch.ones(b, dtype=torch.bool, device=dev) for _ in imgs]
        noise = torch.normal(0.0, 1.0, (b, horizon, 32), generator=torch.Generator().manual_seed(42)).to(dev)
        return imgs, masks, obs["lang_tokens"].to(dev), obs["lang_masks"].to(dev), obs["observation.state"].to(dev), noise

    out = {}
    args_b = inputs(B, 1234)
    model.sample_actions(*args_b)  # warm-up (cuBLAS / cuDNN heuristics, allocator)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        x = model.sample_actions(*args_b)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    assert torch.isfinite(x).all()
    out.update({"value": B / (ms * 1e-3), "unit": "chunks/s", "ms_per_step": ms, "batch": B, "steps": steps})
    if log:
        log(f"gpu eager batch {B}: {ms:.1f} ms/step")
    args_1 = inputs(1, 1234)
    lat = []
    for i in range(3 + b1_iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        model.sample_actions(*args_1)
        b.record()
        torch.cuda.synchronize()
        if i >= 3:
            lat.append(a.elapsed_time(b))
    lat.sort()
    out.update({"p50_b1_ms": lat[len(lat) // 2], "p99_b1_ms": lat[-1], "b1_iters": len(lat),
                "what": "PyTorch eager + torch.autocast(bf16) restatement of the reference's op sequence on this GPU "
                        "(cuBLAS / cuDNN, eager attention, Python Euler loop), inputs resident in HBM; torch "
                        f"{torch.__version__}"})
    del model
    torch.cuda.empty_cache()
    return out


def time_fp32_regime(workload, dev, B=8, steps=2, b1_iters=10, log=None):
    """`--use_bf16 False`, the flag the reference's shipped eval scripts pass (slurms/eval_scripts/simpler/
    ev_pi0_bridge_simpler.sh:99, doc/evaluation.md:152): fp32 weights and autocast(float32), i.e. plain fp32. Here that
    regime is the SIMT fp32 path (csrc/gemm_simt.cu, parity <= 1e-4); this leg times it next to PyTorch eager fp32 (cuBLAS
    SGEMM, TF32 off as torch defaults) on the same GPU, same weights / inputs / noise."""
    from int_act_b200.policy import PI0Policy
    from int_act_b200.weights import Dims, random_state_dict
    from oracle.pi0_eager import Pi0Eager
    from oracle.pi0_oracle import OracleDims

    n_cams, lang_len, horizon = WORKLOADS[workload]
    dims = Dims()
    sd = random_state_dict(dims, device=dev, seed=0)
    policy = PI0Policy(make_config(workload), state_dict=sd, max_batch=B, device=dev)
    policy.to(torch.float32)
    eng = policy._ensure_engine(B)
    policy._state_dict = None

    def timed(fn, n):
        ts = []
        for _ in range(n):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        return ts

    out = {"batch": B, "dtype": "f32"}
    eng._static(B)["x_t"].normal_()
    eng._run_static(B)
    ts = timed(lambda: eng._run_static(B), steps)
    ours_b = sum(ts) / len(ts)
    eng._static(1)["x_t"].normal_()
    eng._run_static(1)
    ts = timed(lambda: eng._run_static(1), b1_iters)
    ours_1 = ts[len(ts) // 2]
    out["b200"] = {"value": B / (ours_b * 1e-3), "unit": "chunks/s", "ms_per_step": ours_b, "p50_b1_ms": ours_1,
                   "path": eng.path_report(B)}
    if log:
        log(f"fp32 regime, this library: batch {B} {ours_b:.1f} ms/step, batch 1 {ours_1:.1f} ms")
    del policy, eng
    torch.cuda.empty_cache()

    model = Pi0Eager(OracleDims(n_cams=n_cams, lang_len=lang_len, horizon=horizon), sd, dev, use_bf16=False)
    del sd
    torch.cuda.empty_cache()

    def inputs(b):
        obs = synthetic_observations(b, n_cams, lang_len, seed=1234, pinned=False)
        imgs = [obs[k].to(dev) for k in obs if k.startswith("observation.images")]
        masks = [torch.ones(b, dtype=torch.bool, device=dev) for _ in imgs]
        noise = torch.normal(0.0, 1.0, (b, horizon, 32), generator=torch.Generator().manual_seed(42)).to(dev)
        return imgs, masks, obs["lang_tokens"].to(dev), obs["lang_masks"].to(dev), obs["observation.state"].to(dev), noise

    ab, a1 = inputs(B), inputs(1)
    model.sample_actions(*ab)
    ts = timed(lambda: model.sample_actions(*ab), steps)
    eager_b = sum(ts) / len(ts)
    model.sample_actions(*a1)
    ts = timed(lambda: model.sample_actions(*a1), b1_iters)
    eager_1 = ts[len(ts) // 2]
    out["eager"] = {"value": B / (eager_b * 1e-3), "unit": "chunks/s", "ms_per_step": eager_b, "p50_b1_ms": eager_1,
                    "what": f"PyTorch eager fp32 (cuBLAS SGEMM, allow_tf32={torch.backends.cuda.matmul.allow_tf32}), torch "
                            f"{torch.__version__}"}
    out["speedup_batch"] = out["b200"]["value"] / out["eager"]["value"]
    out["speedup_b1_latency"] = eager_1 / ours_1
    del model
    torch.cuda.empty_cache()
    return out


def eager_arm(args, rank, local_rank):
    """`--impl eager`: the GPU-eager baseline as its own JSON line (rank 0 only; replicas would time the same thing)."""
    if rank != 0:
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl eager needs a CUDA device")
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    n_cams, lang_len, horizon = WORKLOADS[args.workload]
    sampler = ClockSampler(local_rank)
    sampler.start()
    r = time_gpu_eager(args.workload, args.batch, dev, steps=max(1, args.steps), b1_iters=min(args.latency_iters, 30),
                       log=lambda s: print(s, file=sys.stderr))
    clocks = sampler.result()
    emit({"impl": "eager", "metric": "pi0_action_chunks_per_sec", "value": r["value"], "unit": "chunks/s", "n_gpus": 1,
          "steps": r["steps"], "warmup": 1, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
          "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
          "config": {"workload": f"pi0-{args.workload} infer_action, batch {args.batch}, PyTorch eager + autocast(bf16) on "
                                 "one B200 (reference-equivalent GPU baseline)", "n_cams": n_cams,
                     "prefix_tokens": n_cams * 256 + lang_len, "horizon": horizon, "num_steps": 10},
          "latency_b1": {"p50_ms": r["p50_b1_ms"], "p99_ms": r["p99_b1_ms"], "iters": r["b1_iters"]},
          "gpu_launches": 0, "clocks": clocks, "what": r["what"]})


def reference_arm(args, rank):
    """`--impl reference`: the reference's CPU implementation of the path on the host cores. The reference's own code
    (lerobot PI0Policy) is not in the tree, so this is the oracle port ("kind": "port")."""
    if rank != 0:
        return
    n_cams, lang_len, horizon = WORKLOADS[args.workload]
    # bounded: each step = one observation of the batch; cap the untimed warm-up at 1 step
    times, cores = time_cpu_oracle(args.workload, args.steps, min(args.warmup, 1),
                                   log=lambda s: print(s, file=sys.stderr))
    total = sum(times)
    value = len(times) / total
    sample = (f"each step = infer_action on 1 observation of the {args.batch}-observation batch, fp32, "
              f"{cores} host threads, vocab table cut to 8192 rows")
    line = {
        "impl": "reference", "metric": "pi0_action_chunks_per_sec", "value": value, "unit": "chunks/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": min(args.warmup, 1),
        "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"pi0-{args.workload} infer_action, batch {args.batch}/GPU (reference arm: 1 observation "
                               f"per step on the CPU)", "n_cams": n_cams, "prefix_tokens": n_cams * 256 + lang_len,
                   "horizon": horizon, "num_steps": 10},
        "cpu_baseline": {"value": value, "unit": "chunks/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "chunks/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# -----------------------------------------------------------------------------------------------------------------
# the B200 arm
# -----------------------------------------------------------------------------------------------------------------
def label_flops(label, B, n_cams, lang_len, horizon):
    """Algorithmic FLOPs of ONE launch of the GEMM behind a profiling label."""
    P, S = n_cams * 256 + lang_len, 1 + horizon
    Mg, Me, Ms = B * P, B * S, B * 256
    t = {
        "prefill.gateup_geglu": 2 * Mg * 32768 * 2048, "prefill.down": 2 * Mg * 2048 * 16384,
        "prefill.qkv_rope": 2 * Mg * 2560 * 2048, "prefill.o_proj": 2 * Mg * 2048 * 2048,
        "prefill.scores": 2 * B * P * 8 * P * 256, "prefill.pv": 2 * B * P * 8 * P * 256,
        "prefill.attention": 4 * B * P * 8 * P * 256,
        "siglip.qkv": 2 * Ms * 3456 * 1152, "siglip.out_proj": 2 * Ms * 1152 * 1152,
        "siglip.fc1_gelu": 2 * Ms * 4304 * 1152, "siglip.fc2": 2 * Ms * 4304 * 1152,
        "siglip.scores": 2 * B * 16 * 256 * 256 * 72, "siglip.pv": 2 * B * 16 * 256 * 256 * 72,
        "siglip.patch_embed": 2 * Ms * 1152 * 588, "siglip.projector": 2 * Ms * 2048 * 1152,
        "denoise.gateup_geglu": 2 * Me * 8192 * 1024, "denoise.down": 2 * Me * 1024 * 4096,
        "denoise.qkv_rope": 2 * Me * 2560 * 1024, "denoise.o_proj": 2 * Me * 1024 * 2048,
    }
    return t.get(label)


def b200_arm(args, rank, world, local_rank):
    import torch.distributed as dist

    from int_act_b200 import binding as pb
    from int_act_b200.policy import PI0Policy
    from int_act_b200.weights import Dims, random_state_dict, state_dict_shapes

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device - the sm_100a library is the only implementation (no CPU fallback)")
    pb.lib()  # fail loudly if the extension is missing
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n_cams, lang_len, horizon = WORKLOADS[args.workload]
    B, K, W = args.batch, args.steps, args.warmup
    peaks = measured_peaks()

    # ---- weights: generated once on rank 0, broadcast over NCCL/NVLink ------------------
    dims = Dims()
    sd = random_state_dict(dims, device=dev, seed=0) if rank == 0 else None
    bcast_gbs = None
    if world > 1:
        from int_act_b200.sharding import broadcast_state_dict_blob

        warm = torch.zeros(1 << 20, device=dev)
        for _ in range(2):           # communicator set-up / NVLink channel warm-up outside the timed broadcast
            dist.broadcast(warm, src=0)
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        sd = broadcast_state_dict_blob(sd, state_dict_shapes(dims), torch.bfloat16, dev, src=0)
        torch.cuda.synchronize()
        bcast_s = time.perf_counter() - t0
        bcast_gbs = sum(v.numel() for v in sd.values()) * 2 / bcast_s / 1e9
    else:
        bcast_s = 0.0
    policy = PI0Policy(make_config(args.workload), state_dict=sd, max_batch=B, device=dev)
    policy.to(torch.bfloat16)
    eng = policy._ensure_engine(B)
    del sd
    policy._state_dict = None
    torch.cuda.empty_cache()

    obs = synthetic_observations(B, n_cams, lang_len, seed=1234 + rank, pinned=True)
    h2d = sum(v.numel() * v.element_size() for v in obs.values())
    d2h = B * horizon * 7 * 4

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_max(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- (1) device-resident: inputs already in the engine's static HBM buffers, CUDA-graph replay ----------------
    policy.infer_action(obs)  # stages the inputs, captures the graph
    st = eng._static(B)
    for _ in range(W):
        st["x_t"].normal_()
        eng._run_static(B)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        st["x_t"].normal_()       # fresh N(0,1) noise each step (sample_noise), on the device
        eng._run_static(B)
    e1.record()
    torch.cuda.synchronize()
    clocks = sampler.result()
    barrier()
    ms_total = reduce_max(e0.elapsed_time(e1))
    ms_per_step = ms_total / K
    value = world * B * K / (ms_total / 1e3)
    launches = eng.launches_per_infer.get(B, 0) * K

    # ---- (2) end to end through the public API with pinned host buffers ----------------------------------------
    # N = 1: PI0Policy.infer_action on the host batch. N > 1: sharding.sharded_infer - every rank holds the same
    # [B * N]-observation host batch, runs its contiguous shard and the chunks are gathered over NCCL, so the gathered
    # [B * N, H, 7] array on rank 0's host is INSIDE the timed region.
    if world > 1:
        from int_act_b200.sharding import sharded_infer

        full = synthetic_observations(B * world, n_cams, lang_len, seed=1234, pinned=True)

        def e2e_step():
            return sharded_infer(policy.infer_action, full).cpu()
        h2d = sum(v.numel() * v.element_size() for v in full.values()) // world
        e2e_shape = (B * world, horizon, 7)
    else:
        def e2e_step():
            return policy.infer_action(obs).cpu()   # H2D of the observation, the path, D2H of the chunk
        e2e_shape = (B, horizon, 7)
    for _ in range(max(1, min(W, 2))):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(K):
        out = e2e_step()
    torch.cuda.synchronize()
    e2e_s = reduce_max(time.perf_counter() - t0)
    barrier()
    e2e_value = world * B * K / e2e_s
    assert out.shape == e2e_shape and torch.isfinite(out).all()

    # ---- (2b) the other BASELINE configurations (SURVEY.md 8d "Concrete configs"), all ranks, device-timed ----------------
    # C3 strong split: 64 observations TOTAL (64 / N per GPU); C4 rephrase: 256 total; C5 LIBERO-shaped: 512 total.
    configs = {}
    if not args.no_configs:
        from int_act_b200.engine import Pi0Engine

        plan = [("c3_bridge_strong", "bridge", 64), ("c4_rephrase", "rephrase", 256), ("c5_libero", "libero", 512)]
        for name, wl, total in plan:
            per_gpu = max(1, total // world)
            if name == "c3_bridge_strong" and world == 1 and args.workload == "bridge" and B == 64:
                configs[name] = {"total_batch": 64, "batch_per_gpu": 64, "value": value, "unit": "chunks/s",
                                 "ms_per_step": ms_per_step, "same_as": "headline"}
                continue
            nc, ll, hz = WORKLOADS[wl]
            e = Pi0Engine(eng.packed, n_cams=nc, lang_len=ll, horizon=hz, max_batch=per_gpu)
            ob = synthetic_observations(per_gpu, nc, ll, seed=4321 + rank, pinned=False)
            imgs = [ob[k].to(dev) for k in ob if k.startswith("observation.images")]
            e.sample_actions(imgs, [torch.ones(per_gpu, dtype=torch.bool, device=dev) for _ in imgs],
                             ob["lang_tokens"].to(dev), ob["lang_masks"].to(dev), ob["observation.state"].to(dev))
            stc = e._static(per_gpu)
            for _ in range(2):
                stc["x_t"].normal_()
                e._run_static(per_gpu)
            barrier()
            kc = max(2, min(K, 5))
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(kc):
                stc["x_t"].normal_()
                e._run_static(per_gpu)
            b.record()
            torch.cuda.synchronize()
            ms = reduce_max(a.elapsed_time(b)) / kc
            barrier()
            v = per_gpu * world / (ms * 1e-3)
            tfl = flops_per_chunk(nc, ll, hz) * per_gpu / (ms * 1e-3) / 1e12
            configs[name] = {"total_batch": per_gpu * world, "batch_per_gpu": per_gpu, "value": v, "unit": "chunks/s",
                             "ms_per_step": ms, "steps": kc, "achieved_tflops_per_gpu": tfl, "paths": e.path_report(per_gpu),
                             "frac_of_sustained_peak": tfl / peaks["bf16_tflops_sustained"],
                             "launches_per_infer": e.launches_per_infer.get(per_gpu)}
            del e, stc
            torch.cuda.empty_cache()

    line = {
        "metric": "pi0_action_chunks_per_sec", "value": value, "unit": "chunks/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": f"pi0-{args.workload} infer_action (BASELINE config 3), batch {B} per GPU, random-init "
                               "INTACT-pi0-finetune-bridge architecture", "batch_per_gpu": B, "n_cams": n_cams,
                   "prefix_tokens": n_cams * 256 + lang_len, "horizon": horizon, "num_steps": 10,
                   "parallelism": f"dp{world} (rollout batch sharded, no collective in the forward)",
                   "l2": "inputs larger than L2: every step streams 5.9 GB of weights plus >1 GB of activations "
                         "through the 126 MB L2", "weights_broadcast_s": round(bcast_s, 4),
                   "weights_broadcast_gbs": None if bcast_gbs is None else round(bcast_gbs, 1)},
        "e2e": {"value": e2e_value, "unit": "chunks/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": 1e3 * e2e_s / K},
        "gpu_launches": launches,
        "clocks": clocks,
    }
    if configs:
        line["configs"] = configs
    line["paths"] = {f"batch_{B}": eng.path_report(B), "batch_1": eng.path_report(1)}

    if rank == 0:
        # ---- (3) roofline of the dominant kernel, CUDA-event pair around every launch of one eager step -------------
        # three profiled passes, per-label median (a single pass is at the mercy of the power-cap clock of that moment)
        passes = [eng.profile_step(B) for _ in range(3)]
        prof = {k: (statistics.median(p[k][0] for p in passes), passes[0][k][1]) for k in passes[0]}
        step_ms = sum(v[0] for v in prof.values())
        top = max(prof, key=lambda k: prof[k][0])
        fl = label_flops(top, B, n_cams, lang_len, horizon)
        avg_ms = prof[top][0] / prof[top][1]
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get(f"{top}@B{B}")
        if fl is not None:
            ach = fl / (avg_ms * 1e-3) / 1e12
            line["roofline"] = {"bound": "tensor", "kernel": top, "achieved": ach, "peak": peaks["bf16_tflops_sustained"],
                                "unit": "TFLOP/s", "frac": ach / peaks["bf16_tflops_sustained"], "traffic": traffic,
                                "peak_source": f"{peaks['source']} sustained cuBLAS bf16 (kernel timed inside a step)",
                                "avg_launch_ms": avg_ms, "launches_per_step": prof[top][1],
                                "share_of_step": prof[top][0] / step_ms}
            if peaks.get("bf16_tflops"):
                # the sustained cuBLAS figure was taken at the clock a 4 s back-to-back matmul settles at; a step that
                # interleaves lighter phases runs its GEMMs at a higher clock, so frac can exceed 1: the burst peak bounds it
                line["roofline"]["frac_of_burst_peak"] = ach / peaks["bf16_tflops"]
        line["breakdown_ms"] = {k: round(v[0], 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}
        fpc = flops_per_chunk(n_cams, lang_len, horizon)
        line["whole_path"] = {"tflop_per_chunk": fpc / 1e12,
                              "achieved_tflops_per_gpu": fpc * B / (ms_per_step * 1e-3) / 1e12,
                              "frac_of_sustained_peak": fpc * B / (ms_per_step * 1e-3) / 1e12 / peaks["bf16_tflops_sustained"]}

        # ---- (3b) the reference-facing call: LeRobotPolicyWrapper.select_action on raw uint8 frames ------------------
        # (GPU adapter -> policy -> GPU post-processing -> float64 env actions on the host), what the policy server runs
        if n_cams == 1:
            import numpy as np

            from int_act_b200 import adapters as ad
            from int_act_b200.policy import SyntheticTokenizer
            from int_act_b200.serving import LeRobotPolicyWrapper

            stats = {"proprio": {"p01": [-0.5] * 7, "p99": [0.5] * 7}, "action": {"p01": [-1.0] * 7, "p99": [1.0] * 7}}
            adapter = ad.BridgeSimplerBatchAdapter(ad.adapter_config("synthetic", use_bf16=True), dataset_statistics=stats,
                                                   device=dev)
            if policy._tokenizer is None:
                policy._tokenizer = SyntheticTokenizer()
            wrapper = LeRobotPolicyWrapper(policy, adapter, action_step=horizon)
            rng = np.random.default_rng(1234)
            eef = rng.normal(size=(B, 8)).astype(np.float32)
            raw = {"observation.images.top": rng.integers(0, 256, (B, 480, 640, 3), dtype=np.uint8),
                   "observation.state": eef, "task": ["put the carrot on the plate"] * B}
            for _ in range(2):
                env_actions = wrapper.select_action(raw)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(K):
                env_actions = wrapper.select_action(raw)
            wr_s = time.perf_counter() - t0
            assert env_actions.shape == (B, horizon, 7) and env_actions.dtype == np.float64
            line["e2e_wrapper"] = {"value": B * K / wr_s, "unit": "chunks/s", "ms_per_step": 1e3 * wr_s / K,
                                   "h2d_bytes_per_step": int(raw["observation.images.top"].nbytes + eef.nbytes),
                                   "d2h_bytes_per_step": int(env_actions.nbytes),
                                   "what": "LeRobotPolicyWrapper.select_action on pageable uint8 480x640 frames + eef "
                                           "poses + task strings (rank 0): GPU adapters, tokeniser, engine, GPU "
                                           "post-processing, float64 actions back on the host"}

        # ---- (4) batch-1 latency (BASELINE config 2): p50 / p99 and the HBM-roofline fraction ---------------------
        obs1 = {k: v[:1].clone().pin_memory() for k, v in obs.items()}
        policy.infer_action(obs1)
        st1 = eng._static(1)
        lat = []
        for i in range(10 + args.latency_iters):
            st1["x_t"].normal_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            eng._run_static(1)
            b.record()
            torch.cuda.synchronize()
            if i >= 10:
                lat.append(a.elapsed_time(b))
        lat.sort()
        prof1 = eng.profile_step(1)
        line["breakdown_b1_ms"] = {k: round(v[0], 4) for k, v in sorted(prof1.items(), key=lambda kv: -kv[1][0])}
        e2e_lat = []
        for i in range(5 + args.latency_iters):
            t0 = time.perf_counter()
            policy.infer_action(obs1).cpu()
            if i >= 5:
                e2e_lat.append(1e3 * (time.perf_counter() - t0))
        e2e_lat.sort()
        p50 = lat[len(lat) // 2]
        by = bytes_per_chunk_b1(n_cams, lang_len, horizon)
        line["latency_b1"] = {
            "p50_ms": p50, "p99_ms": lat[min(len(lat) - 1, int(len(lat) * 0.99))], "iters": len(lat),
            "e2e_p50_ms": e2e_lat[len(e2e_lat) // 2], "e2e_p99_ms": e2e_lat[min(len(e2e_lat) - 1, int(len(e2e_lat) * 0.99))],
            "gpu_launches_per_infer": eng.launches_per_infer.get(1),
            "roofline": {"bound": "hbm", "achieved": by / (p50 * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": by / (p50 * 1e-3) / 1e9 / peaks["hbm_gbs"], "algorithmic_bytes": by,
                         "peak_source": peaks["source"]}}
        rl_ms, rl_ph = roofline_ms_b1(n_cams, lang_len, horizon, peaks)
        line["latency_b1"]["roofline_per_phase"] = {
            "bound": "per phase, the slower of FLOPs at the burst bf16 peak and weight + KV bytes at the copy bandwidth",
            "roofline_ms": rl_ms, "frac": rl_ms / p50, "phases": {k: {a: round(b, 4) for a, b in v.items()}
                                                                 for k, v in rl_ph.items()}}

        # ---- (4b) the reference-equivalent GPU number: PyTorch eager + autocast(bf16) on this same B200 -------------------
        del policy, eng, st, st1
        torch.cuda.empty_cache()
        if not args.no_gpu_eager:
            ge = time_gpu_eager(args.workload, B, dev, steps=3, b1_iters=20)
            ge["speedup_batch"] = value / world / ge["value"]
            ge["speedup_b1_latency"] = ge["p50_b1_ms"] / p50
            line["gpu_eager"] = ge
        if args.fp32_regime and world == 1:
            line["fp32_regime"] = time_fp32_regime(args.workload, dev, B=args.fp32_batch)

        # ---- (5) CPU baseline on the host cores (N=1 only) --------------------------------------------------------
        if world == 1 and not args.no_cpu_baseline:
            times, cores = time_cpu_oracle(args.workload, 2, 1)
            v = len(times) / sum(times)
            line["cpu_baseline"] = {"value": v, "unit": "chunks/s", "cores": cores, "kind": "port",
                                    "sample": f"fp32 CPU oracle, {len(times)} timed infer_action calls on 1 observation "
                                              f"after 1 warm-up ({sum(times):.1f} s), vocab table cut to 8192 rows"}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "eager"])
    ap.add_argument("--batch", type=int, default=64, help="observations per GPU per step")
    ap.add_argument("--workload", default="bridge", choices=sorted(WORKLOADS))
    ap.add_argument("--latency-iters", type=int, default=100)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-eager", action="store_true")
    ap.add_argument("--fp32-regime", action="store_true",
                    help="also time the --use_bf16 False regime (SIMT fp32 path) next to PyTorch eager fp32")
    ap.add_argument("--fp32-batch", type=int, default=8, help="rollout batch of the --fp32-regime leg")
    ap.add_argument("--no-configs", action="store_true", help="skip the C3-strong / C4 / C5 side measurements")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    if args.impl == "reference":
        reference_arm(args, rank)
    elif args.impl == "eager":
        eager_arm(args, rank, local_rank)
    else:
        b200_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
