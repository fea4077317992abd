"""Read-batch sharding helpers: reads are independent, so N GPUs each take a contiguous range of
reads with no collective on the data path; only the results are gathered, in submission order.

* ``merge_results``  -- concatenate per-shard results (offsets rebased) on one host
* ``gather_results`` -- the same across ``torch.distributed`` ranks (one process per GPU; gloo or nccl group),
  delivered on rank ``dst``
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np

from .api import BatchResult, partition_reads  # noqa: F401  (re-exported)


def merge_results(parts: List[BatchResult]) -> BatchResult:
    """Concatenate shard results in shard order; u_off / b_off are rebased onto the merged arrays."""
    if not parts:
        z = np.zeros(1, np.int64)
        return BatchResult(np.zeros(0, np.int32), z, z.copy(), np.zeros(0, np.uint64), np.zeros((0, 2), np.uint64))
    n_u = np.concatenate([p.n_u for p in parts])
    u = np.concatenate([p.u for p in parts])
    b = np.concatenate([p.b.reshape(-1, 2) for p in parts], axis=0)
    u_off, b_off, ub, bb = [], [], 0, 0
    for p in parts:
        u_off.append(p.u_off[:-1] + ub)
        b_off.append(p.b_off[:-1] + bb)
        ub += int(p.u_off[-1])
        bb += int(p.b_off[-1])
    u_off.append(np.array([ub], np.int64))
    b_off.append(np.array([bb], np.int64))
    return BatchResult(n_u, np.concatenate(u_off), np.concatenate(b_off), u, b,
                       n_units=sum(p.n_units for p in parts), n_sub_batches=sum(p.n_sub_batches for p in parts),
                       n_visited=sum(p.n_visited for p in parts), n_lanes=sum(p.n_lanes for p in parts))


def shard_of(offs, rank: int, world: int):
    """-> (r0, r1): the contiguous read range rank ``rank`` of ``world`` owns."""
    cuts = partition_reads(offs, world)
    return int(cuts[rank]), int(cuts[rank + 1])


def gather_results(local: BatchResult, dst: int = 0, group=None) -> Optional[BatchResult]:
    """Gather every rank's shard result on rank ``dst`` (in rank order = read order).  Host-side only: tensors travel
    over the process group's CPU path (gloo) or are staged through the device for nccl."""
    import torch
    import torch.distributed as dist

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")

    def allgather_var(arr: np.ndarray, dtype):
        flat = torch.from_numpy(np.ascontiguousarray(arr).view(dtype).reshape(-1).copy()).to(dev)
        sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
        dist.all_gather(sizes, torch.tensor([flat.numel()], dtype=torch.int64, device=dev), group=group)
        sizes = [int(s.item()) for s in sizes]
        m = max(max(sizes), 1)
        pad = torch.zeros(m, dtype=flat.dtype, device=dev)
        pad[:flat.numel()] = flat
        bufs = [torch.zeros(m, dtype=flat.dtype, device=dev) for _ in range(world)]
        dist.all_gather(bufs, pad, group=group)
        return [bufs[k][:sizes[k]].cpu().numpy() for k in range(world)]

    n_u = allgather_var(local.n_u, np.int32)
    u_off = allgather_var(local.u_off, np.int64)
    b_off = allgather_var(local.b_off, np.int64)
    u = allgather_var(local.u.view(np.int64), np.int64)
    b = allgather_var(local.b.reshape(-1).view(np.int64), np.int64)
    meta = allgather_var(np.array([local.n_units, local.n_sub_batches, local.n_visited], np.int64), np.int64)
    if rank != dst:
        return None
    parts = [BatchResult(n_u[k], u_off[k], b_off[k], u[k].view(np.uint64), b[k].view(np.uint64).reshape(-1, 2),
                         int(meta[k][0]), int(meta[k][1]), int(meta[k][2])) for k in range(world)]
    return merge_results(parts)
