"""Rollout-batch sharding across the GPUs of one box.

The path shards by observation: `sample_actions` has no cross-sample op, and the reference itself scales out with
independent server processes (/root/reference/doc/evaluation.md:119-135). Here: one process per GPU, the batch is
split contiguously ON THE HOST before H2D (observations never cross NVLink), every rank runs its slice through its
own engine, and the action chunks (f32 [B/G, H, action_dim], <= 29 KB per rank) are gathered. Weights are loaded or
generated on rank 0 and broadcast once. There is no collective inside the forward pass.

Works on any `torch.distributed` backend: NCCL on the GPUs, gloo in the CPU tests of the host logic.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous [lo, hi) of `n` observations owned by `rank`; the first n % world ranks take one extra."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"rank {rank} outside world {world}")
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_batch(batch: Dict[str, object], world: int, rank: int) -> Dict[str, object]:
    """Slice every per-observation entry of a policy batch (tensors / arrays along dim 0, `task` lists)."""
    n = None
    for v in batch.values():
        if isinstance(v, torch.Tensor) and v.ndim >= 1:
            n = v.shape[0]
            break
    if n is None:
        raise ValueError("batch holds no tensor to take the batch size from")
    lo, hi = shard_bounds(n, world, rank)
    out = {}
    for k, v in batch.items():
        if isinstance(v, torch.Tensor) and v.ndim >= 1 and v.shape[0] == n:
            out[k] = v[lo:hi]
        elif isinstance(v, (list, tuple)) and len(v) == n:
            out[k] = list(v[lo:hi])
        else:
            out[k] = v  # scalars / a single task string shared by the whole batch
    return out


def broadcast_state_dict(sd: Optional[Dict[str, torch.Tensor]], shapes: Dict[str, Tuple[int, ...]], dtype, device,
                         src: int = 0, group=None) -> Dict[str, torch.Tensor]:
    """Rank `src` holds `sd`; every other rank allocates from `shapes` and receives. One broadcast per tensor, in
    sorted key order on every rank (NCCL over NVLink on the GPU box: ~6.5 GB once per model load)."""
    rank = dist.get_rank(group)
    out = {}
    for k in sorted(shapes):
        if rank == src:
            t = sd[k].to(device=device, dtype=dtype).contiguous()
        else:
            t = torch.empty(shapes[k], dtype=dtype, device=device)
        dist.broadcast(t, src=src, group=group)
        out[k] = t
    return out


def broadcast_state_dict_blob(sd: Optional[Dict[str, torch.Tensor]], shapes: Dict[str, Tuple[int, ...]], dtype, device,
                              src: int = 0, group=None, pieces: int = 1) -> Dict[str, torch.Tensor]:
    """The whole state dict as ONE contiguous blob and one collective ("ncclBroadcast of the packed weight
    blob ... one message"): rank `src` flattens its tensors into the blob, every rank receives it and re-views it per key
    (sorted key order). Round 1 issued ~900 per-tensor broadcasts (6 GB/s at N = 8, dominated by per-call latency and
    lazy connection set-up); one ~6.9 GB message runs at link speed. `pieces` > 1 splits the blob into that many equal
    broadcasts (bounded staging). Call after the communicator is warm (any small collective) when timing it."""
    rank = dist.get_rank(group)
    keys = sorted(shapes)
    numels = [int(torch.Size(shapes[k]).numel()) for k in keys]
    total = sum(numels)
    blob = torch.empty(total, dtype=dtype, device=device)
    if rank == src:
        o = 0
        for k, n in zip(keys, numels):
            if tuple(sd[k].shape) != tuple(shapes[k]):
                raise ValueError(f"{k}: shape {tuple(sd[k].shape)} != expected {tuple(shapes[k])}")
            blob[o:o + n].copy_(sd[k].reshape(-1))
            o += n
    pieces = max(1, int(pieces))
    step = (total + pieces - 1) // pieces
    for i in range(pieces):
        part = blob[i * step: min(total, (i + 1) * step)]
        if part.numel():
            dist.broadcast(part, src=src, group=group)
    out, o = {}, 0
    for k, n in zip(keys, numels):
        out[k] = blob[o:o + n].view(shapes[k])
        o += n
    return out


def gather_chunks(local: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """all_gather of the per-rank action chunks [n_local, H, A] -> [n_total, H, A] in observation order (every
    rank gets the full result). Ragged shards are padded to the largest shard for the collective."""
    world = dist.get_world_size(group)
    sizes = [shard_bounds(n_total, world, r) for r in range(world)]
    cap = max(hi - lo for lo, hi in sizes)
    if local.shape[0] != sizes[dist.get_rank(group)][1] - sizes[dist.get_rank(group)][0]:
        raise ValueError(f"rank {dist.get_rank(group)} holds {local.shape[0]} chunks, expected "
                         f"{sizes[dist.get_rank(group)]}")
    pad = torch.zeros((cap,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    parts: List[torch.Tensor] = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[: hi - lo] for p, (lo, hi) in zip(parts, sizes)], dim=0)


def sharded_infer(infer_fn: Callable[[Dict[str, object]], torch.Tensor], batch: Dict[str, object],
                  group=None) -> torch.Tensor:
    """Every rank calls this with the SAME full batch (host memory); each runs `infer_fn` on its shard only and the
    chunks are gathered. `infer_fn` is `PI0Policy.infer_action` on the GPU box."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n = next(v.shape[0] for v in batch.values() if isinstance(v, torch.Tensor) and v.ndim >= 1)
    local = shard_batch(batch, world, rank)
    lo, hi = shard_bounds(n, world, rank)
    if hi > lo:
        out = infer_fn(local)
    else:  # more ranks than observations: this rank idles but still joins the collective
        out = None
    shape_src = torch.zeros(2, dtype=torch.int64, device=out.device if out is not None else "cpu")
    if out is not None:
        shape_src[0], shape_src[1] = out.shape[1], out.shape[2]
    if dist.get_backend(group) == "nccl" and shape_src.device.type != "cuda":
        shape_src = shape_src.cuda()
    dist.all_reduce(shape_src, op=dist.ReduceOp.MAX, group=group)
    if out is None:
        out = torch.zeros((0, int(shape_src[0]), int(shape_src[1])), dtype=torch.float32, device=shape_src.device)
    return gather_chunks(out, n, group)
