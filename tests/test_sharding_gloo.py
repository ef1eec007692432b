"""World-size-2 `gloo` test (CPU) of the multi-GPU host logic: contiguous rollout-batch sharding, weight broadcast
from rank 0, gather of the action chunks in observation order. The per-shard "inference" is a
deterministic stand-in: the CUDA engine has no CPU fallback and is covered by the gpu tests."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from int_act_b200.sharding import (broadcast_state_dict, broadcast_state_dict_blob, shard_batch, shard_bounds,  # noqa: E402
                                   sharded_infer)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_infer(batch):
    # chunk[b, h, a] depends only on observation b's own inputs -> order / ownership mistakes are visible
    s = batch["observation.state"].float()
    base = s.sum(dim=1) + batch["lang_tokens"].float().sum(dim=1) + float(len(batch["task"]) * 0)
    h = torch.arange(4, dtype=torch.float32)[None, :, None]
    a = torch.arange(7, dtype=torch.float32)[None, None, :]
    return base[:, None, None] + 10.0 * h + a


def _worker(rank, world, port, n_obs, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(7)
        batch = {
            "observation.state": torch.randn((n_obs, 7), generator=g),
            "lang_tokens": torch.randint(0, 1000, (n_obs, 72), generator=g),
            "task": [f"task {i}" for i in range(n_obs)],
            "shared": 3,
        }
        shapes = {"a.weight": (4, 3), "b.bias": (5,)}
        sd = {k: torch.full(s, float(i + 1)) for i, (k, s) in enumerate(sorted(shapes.items()))} if rank == 0 else None
        got = broadcast_state_dict(sd, shapes, torch.float32, "cpu")
        ok_bcast = all(torch.equal(got[k], torch.full(shapes[k], float(i + 1))) for i, k in enumerate(sorted(shapes)))
        for pieces in (1, 3):   # the one-blob broadcast (what bench.py uses at N > 1)
            blob = broadcast_state_dict_blob(sd, shapes, torch.float32, "cpu", pieces=pieces)
            ok_bcast = ok_bcast and all(torch.equal(blob[k], torch.full(shapes[k], float(i + 1)))
                                        for i, k in enumerate(sorted(shapes))) and blob["a.weight"].shape == (4, 3)
        out = sharded_infer(_fake_infer, batch)
        full = _fake_infer(batch)
        q.put((rank, ok_bcast, torch.equal(out, full), tuple(out.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_obs", [8, 5, 1])
def test_shard_gather_world2_gloo(n_obs):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_obs, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok_b, ok_g, shape in res:
        assert ok_b, f"rank {rank}: broadcast mismatch"
        assert ok_g, f"rank {rank}: gathered chunks differ from the unsharded result"
        assert shape == (n_obs, 4, 7)


def test_shard_bounds_cover_and_are_contiguous():
    for n in (0, 1, 5, 64, 65, 511):
        for world in (1, 2, 3, 8):
            spans = [shard_bounds(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(4, 2, 2)


def test_shard_batch_slices_lists_and_keeps_shared():
    b = {"x": torch.arange(10).reshape(5, 2), "task": list("abcde"), "k": "same"}
    s = shard_batch(b, 2, 1)
    assert torch.equal(s["x"], b["x"][3:]) and s["task"] == ["d", "e"] and s["k"] == "same"
