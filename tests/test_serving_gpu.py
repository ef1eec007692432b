"""End to end through the reference-facing surface on the GPU: websocket client -> server -> LeRobotPolicyWrapper ->
GPU batch adapter -> PI0Policy (C ABI) -> GPU post-processing -> float64 actions on the wire, checked against the same
chain restated on the CPU (adapter oracle -> Pi0 oracle -> adapter oracle). Reduced depth (2 + 2 layers), full widths."""
import os
import threading

import numpy as np
import pytest
import torch

from oracle import adapters_oracle as ao
from pi0_helpers import oracle_for
from test_serving_cpu import ServerThread

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _build(cuda, max_batch):
    from int_act_b200 import adapters as ad
    from int_act_b200.config import bridge_config
    from int_act_b200.policy import PI0Policy, SyntheticTokenizer
    from int_act_b200.serving import LeRobotPolicyWrapper
    from int_act_b200.weights import Dims, random_state_dict

    d = np.load(os.path.join(G, "adapter_batch.npz"))
    stats = {"proprio": {"p01": d["p01"].tolist(), "p99": d["p99"].tolist()},
             "action": {"p01": d["a01"].tolist(), "p99": d["a99"].tolist()}}
    dims = Dims(n_siglip_layers=2, n_gemma_layers=2, vocab=2048)
    sd = random_state_dict(dims, device=cuda, seed=0)
    tok = SyntheticTokenizer(dims.vocab)
    pol = PI0Policy(bridge_config(), state_dict=sd, dims=dims, max_batch=max_batch, tokenizer=tok, device=cuda)
    pol.to(torch.bfloat16).eval()
    adapter = ad.BridgeSimplerBatchAdapter(ad.adapter_config("unused", use_bf16=True), dataset_statistics=stats)
    return LeRobotPolicyWrapper(pol, adapter, action_step=4), sd, dims, stats, tok


def _cpu_chain(sd, dims, stats, tok, frames, eef, tasks, noise):
    imgs, state = ao.bridge_batch_preprocess(frames, eef, stats)
    state = torch.as_tensor(state, dtype=torch.bfloat16).float()                    # adapter dtype rounding
    ids, mask = tok([t + "\n" for t in tasks], 72)
    orc = oracle_for(sd, dims, 1, 72, 4, "bf16")
    B = len(tasks)
    act = orc.infer_action([torch.from_numpy(imgs)], [torch.ones(B, dtype=torch.bool)], ids, mask, state, noise)
    return ao.bridge_postprocess(act.float().numpy(), stats)


def test_client_server_engine_roundtrip_matches_cpu_chain(cuda):
    from int_act_b200.serving import WebsocketPolicyClient

    wrapper, sd, dims, stats, tok = _build(cuda, max_batch=2)
    rng = np.random.default_rng(0)
    frames = rng.integers(0, 256, (2, 480, 640, 3), dtype=np.uint8)
    eef = np.load(os.path.join(G, "adapter_batch.npz"))["eef"][:2].astype(np.float32)
    tasks = ["put the carrot on the plate", "stack the green block on the yellow block"]
    noise = torch.normal(0.0, 1.0, (2, 4, 32), generator=torch.Generator().manual_seed(42))
    want = _cpu_chain(sd, dims, stats, tok, frames, eef, tasks, noise)

    # pin the noise the policy draws (the reference draws it from the global generator inside sample_actions)
    eng_sample = wrapper.model._ensure_engine(2).sample_actions
    wrapper.model._ensure_engine(2).sample_actions = lambda *a, **k: eng_sample(*a[:5], noise.to(cuda)[: a[4].shape[0]])
    with ServerThread(wrapper) as srv:
        c = WebsocketPolicyClient("127.0.0.1", srv.bound_port)
        assert c.reset() == {"status": "reset"}
        got = c.infer({"observation.images.top": frames, "observation.state": eef, "task": tasks})
        c.close()
    assert got.dtype == np.float64 and got.shape == (2, 4, 7)
    # translation / rotation dims: bf16 tolerance of BASELINE.json (1e-2 on the normalised action) scaled by the
    # de-normalisation range (<= 1 here); gripper is binarised
    assert np.abs(got[..., :3] - want[..., :3]).max() <= 1e-2
    assert np.abs(got[..., 3:6] - want[..., 3:6]).max() <= 2e-2
    assert np.array_equal(got[..., 6], want[..., 6])


def test_checkpoint_hot_swap_through_the_server(cuda, tmp_path):
    """`switch_model` over the wire reloads a lerobot-layout `model.safetensors` (from_pretrained), and
    the next inference runs on the new weights (policy_wrapper.py:61-67,141-170; websocket_policy_server.py:66-71)."""
    import time

    from int_act_b200 import adapters as ad
    from int_act_b200.config import bridge_config
    from int_act_b200.policy import PI0Policy, SyntheticTokenizer
    from int_act_b200.serving import LeRobotPolicyWrapper, WebsocketPolicyClient
    from int_act_b200.weights import Dims, random_state_dict, save_safetensors

    d = np.load(os.path.join(G, "adapter_batch.npz"))
    stats = {"proprio": {"p01": d["p01"].tolist(), "p99": d["p99"].tolist()},
             "action": {"p01": d["a01"].tolist(), "p99": d["a99"].tolist()}}
    dims = Dims(n_siglip_layers=1, n_gemma_layers=2, vocab=1024)
    paths = []
    for seed in (0, 1):
        p = str(tmp_path / f"step_{seed}")
        save_safetensors({k: v.cpu() for k, v in random_state_dict(dims, device=cuda, seed=seed).items()}, p)
        paths.append(p)
    tok = SyntheticTokenizer(dims.vocab)

    def load(path):
        pol = PI0Policy.from_pretrained(pretrained_name_or_path=path, config=bridge_config(), tokenizer=tok, device=cuda)
        pol.model.paligemma_with_expert.gemma_expert.lm_head = None          # what the reference wrapper does
        pol.model.paligemma_with_expert.paligemma.language_model.lm_head = None
        return pol.to(torch.bfloat16).to(cuda).eval()

    adapter = ad.BridgeSimplerBatchAdapter(ad.adapter_config("unused", use_bf16=True), dataset_statistics=stats)
    wrapper = LeRobotPolicyWrapper(load(paths[0]), adapter, action_step=4, model_loader=load)
    rng = np.random.default_rng(3)
    obs = {"observation.images.top": rng.integers(0, 256, (1, 480, 640, 3), dtype=np.uint8),
           "observation.state": d["eef"][:1].astype(np.float32), "task": ["pick up the spoon"]}
    with ServerThread(wrapper) as srv:
        c = WebsocketPolicyClient("127.0.0.1", srv.bound_port)
        torch.manual_seed(7)
        a0 = c.infer(obs)
        t0 = time.perf_counter()
        assert c.switch_model(paths[1]) == {"status": "model switched"}
        swap_s = time.perf_counter() - t0
        torch.manual_seed(7)
        a1 = c.infer(obs)
        assert c.switch_model(paths[0]) == {"status": "model switched"}
        torch.manual_seed(7)
        a0b = c.infer(obs)
        c.close()
    print(f"\n[hot swap] switch_model round trip (reduced-depth checkpoint): {1e3 * swap_s:.1f} ms")
    assert a0.shape == (4, 7) and a0.dtype == np.float64
    assert np.array_equal(a0, a0b)                 # same checkpoint + same seed => same actions (deterministic path)
    assert np.abs(a0 - a1).max() > 1e-3            # the swapped-in weights are really in use


def test_replicas_behind_one_port_on_the_gpu_and_worker_thread_device_pinning(cuda):
    """Two engine replicas behind one websocket port (on two GPUs when the box has them, else both on cuda:0), driven from
    the server's worker threads: exercises the per-device kernel attributes / device pinning (a worker thread's current
    device is cuda:0 whatever device its engine lives on), request coalescing with pre-captured graphs and the latency
    histogram. Every client must get the action of ITS observation whichever replica served it."""
    from int_act_b200.serving import WebsocketPolicyClient, WebsocketPolicyServer

    n_dev = torch.cuda.device_count()
    devs = [torch.device("cuda", 0), torch.device("cuda", 1 if n_dev > 1 else 0)]
    wrappers = []
    for d in devs:
        w, sd, dims, stats, tok = _build(d, max_batch=4)
        w.env_adapter = type(w.env_adapter)(__import__("int_act_b200.adapters", fromlist=["x"]).adapter_config(
            "unused", use_bf16=True), dataset_statistics=stats, device=d)
        w.max_batch = 4
        w.warmup(4)                      # graphs for batch 1..4 captured before the first request
        assert sorted(w.model._engine._graphs) == [1, 2, 3, 4]
        wrappers.append(w)
    rng = np.random.default_rng(1)
    n = 8
    frames = rng.integers(0, 256, (n, 480, 640, 3), dtype=np.uint8)
    eef = np.load(os.path.join(G, "adapter_batch.npz"))["eef"]
    eef = np.concatenate([eef] * (n // len(eef) + 1))[:n].astype(np.float32)
    noise = torch.normal(0.0, 1.0, (1, 4, 32), generator=torch.Generator().manual_seed(42))
    for w in wrappers:   # same noise row for every observation -> the result depends only on the observation
        eng = w.model._engine
        orig = eng.sample_actions
        eng.sample_actions = lambda *a, orig=orig, eng=eng, **k: orig(*a[:5], noise.to(eng.device).expand(a[4].shape[0], 4, 32))
    single = [wrappers[0].select_action({"observation.images.top": frames[i:i + 1], "observation.state": eef[i:i + 1],
                                         "task": ["put the carrot on the plate"]}) for i in range(n)]
    srv_obj = ServerThread.__new__(ServerThread)
    srv_obj.server = WebsocketPolicyServer(wrappers[0], host="127.0.0.1", port=0, metadata={}, max_batch=4,
                                           batch_window_ms=20.0, replicas=wrappers[1:])
    srv_obj.ready = threading.Event()
    srv_obj.thread = threading.Thread(target=srv_obj._main, daemon=True)
    with srv_obj as srv:
        clients = [WebsocketPolicyClient("127.0.0.1", srv.bound_port) for _ in range(n)]
        out = [None] * n

        def run(i):
            for _ in range(3):
                out[i] = clients[i].infer({"observation.images.top": frames[i:i + 1], "observation.state": eef[i:i + 1],
                                           "task": ["put the carrot on the plate"]})
        ts = [threading.Thread(target=run, args=(i,)) for i in range(n)]
        [t.start() for t in ts]
        [t.join(120) for t in ts]
        stats_srv, lat = dict(srv.stats), srv.latency_ms()
        [c.close() for c in clients]
    for i in range(n):
        assert out[i] is not None and out[i].shape == single[i].shape
        assert np.abs(out[i][..., :6] - single[i][..., :6]).max() <= 2e-2, i      # batch composition changes bf16 rounding paths
    assert stats_srv["requests"] == 3 * n and stats_srv["failed_requests"] == 0
    assert sum(1 for b in stats_srv["per_replica_batches"] if b > 0) == 2, stats_srv
    assert lat["n"] == 3 * n and lat["p99"] >= lat["p50"] > 0
    print(f"\n[2 replicas on {n_dev} GPU(s)] {stats_srv}; latency ms {lat}")
