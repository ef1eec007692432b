"""Requests/s through the websocket policy server (row f2): N engine replicas behind ONE port (one per visible GPU),
C client threads each sending single-observation requests (uint8 480x640x3 frame + eef pose + task string, 0.92 MB on the
wire, what SimplerEnv sends) back to back for T seconds. Prints one JSON line: requests/s, server-side p50/p99 request
latency, batches formed. Run on the GPU box:  python tools/server_bench.py [--gpus N] [--clients C] [--max-batch M] [--seconds T]"""
import argparse
import asyncio
import json
import os
import sys
import threading
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from int_act_b200 import adapters as ad  # noqa: E402
from int_act_b200.config import bridge_config  # noqa: E402
from int_act_b200.policy import PI0Policy, SyntheticTokenizer  # noqa: E402
from int_act_b200.serving import LeRobotPolicyWrapper, WebsocketPolicyClient, WebsocketPolicyServer  # noqa: E402
from int_act_b200.weights import Dims, random_state_dict  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=torch.cuda.device_count())
    ap.add_argument("--clients", type=int, default=64)
    ap.add_argument("--max-batch", type=int, default=32)
    ap.add_argument("--workers", type=int, default=1, help="server worker threads per replica")
    ap.add_argument("--window-ms", type=float, default=2.0)
    ap.add_argument("--seconds", type=float, default=10.0)
    a = ap.parse_args()
    stats = {"proprio": {"p01": [-0.5] * 7, "p99": [0.5] * 7}, "action": {"p01": [-1.0] * 7, "p99": [1.0] * 7}}
    dims = Dims(vocab=8192)
    wrappers = []
    for g in range(a.gpus):
        dev = torch.device("cuda", g)
        sd = random_state_dict(dims, device=dev, seed=0)
        pol = PI0Policy(bridge_config(), state_dict=sd, dims=dims, max_batch=a.max_batch, tokenizer=SyntheticTokenizer(dims.vocab),
                        device=dev)
        pol.to(torch.bfloat16).eval()
        adapter = ad.BridgeSimplerBatchAdapter(ad.adapter_config("unused", use_bf16=True), dataset_statistics=stats, device=dev)
        w = LeRobotPolicyWrapper(pol, adapter, action_step=4, max_batch=a.max_batch)
        w.warmup(a.max_batch)
        pol._state_dict = None
        del sd
        wrappers.append(w)
    server = WebsocketPolicyServer(wrappers[0], host="127.0.0.1", port=0, metadata={}, max_batch=a.max_batch, workers_per_replica=a.workers,
                                   batch_window_ms=a.window_ms, replicas=wrappers[1:])
    ready, holder = threading.Event(), {}

    def serve():
        async def go():
            holder["loop"] = asyncio.get_running_loop()
            holder["stop"] = asyncio.Event()
            r = asyncio.Event()
            task = asyncio.create_task(server.run(ready=r, stop=holder["stop"]))
            await r.wait()
            ready.set()
            await task
        asyncio.run(go())

    th = threading.Thread(target=serve, daemon=True)
    th.start()
    assert ready.wait(30)
    # clients live in separate PROCESSES (their msgpack packing must not share the server's GIL)
    import multiprocessing as mp

    ctx = mp.get_context("spawn")
    n_proc = min(8, a.clients)
    per = [a.clients // n_proc + (1 if i < a.clients % n_proc else 0) for i in range(n_proc)]
    q = ctx.Queue()
    go, t_end = ctx.Event(), ctx.Value("d", 0.0)
    procs = [ctx.Process(target=client_proc, args=(server.bound_port, per[i], i, q, go, t_end)) for i in range(n_proc)]
    [p.start() for p in procs]
    for _ in range(n_proc):
        assert q.get(timeout=300) == "ready"
    server._lat_ms.clear()
    b0 = server.stats["batches"]
    t0 = time.time()
    t_end.value = t0 + a.seconds
    go.set()
    total = sum(q.get(timeout=a.seconds + 300) for _ in range(n_proc))
    dt = time.time() - t0
    [p.join(30) for p in procs]
    holder["loop"].call_soon_threadsafe(holder["stop"].set)
    print(json.dumps({"gpus": a.gpus, "clients": a.clients, "client_processes": n_proc, "max_batch": a.max_batch,
                      "window_ms": a.window_ms, "requests_per_s": total / dt, "seconds": dt,
                      "latency_ms": server.latency_ms(), "batches": server.stats["batches"] - b0,
                      "max_batch_seen": server.stats["max_batch_seen"],
                      "per_replica_batches": server.stats["per_replica_batches"], "wire_bytes_per_request": 921600 + 32}))


def client_proc(port, n_threads, seed, q, go, t_end):
    rng = np.random.default_rng(seed)
    obs = [{"observation.images.top": rng.integers(0, 256, (1, 480, 640, 3), dtype=np.uint8),
            "observation.state": rng.normal(size=(1, 8)).astype(np.float32), "task": ["put the carrot on the plate"]}
           for _ in range(n_threads)]
    clients = [WebsocketPolicyClient("127.0.0.1", port) for _ in range(n_threads)]
    for c, o in zip(clients, obs):
        c.infer(o)   # connection + first request outside the timed window
    counts = [0] * n_threads
    q.put("ready")
    go.wait()

    def run(i):
        while time.time() < t_end.value:
            out = clients[i].infer(obs[i])
            assert out.shape == (4, 7)
            counts[i] += 1

    ts = [threading.Thread(target=run, args=(i,)) for i in range(n_threads)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    [c.close() for c in clients]
    q.put(sum(counts))


if __name__ == "__main__":
    main()
