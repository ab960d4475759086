"""Multi-GPU plumbing: one process per GPU, ``torch.distributed`` (NCCL on GPUs, gloo in the CPU tests).

Two ways the path shards:
* replicates -- independent units, the reference's own scheme (``worker_filter``,
  /root/reference/src/cnmf/cnmf.py:53-54, 876-880): no collective at all;
* cells (online mode, very large N) -- each rank owns a contiguous block of rows of X and of H; the only
  exchange is the sum of the per-chunk sufficient statistics A = h^T h (R x Kp x Kp) and B = h^T x
  (R*Kp x G) and of the objective's partial terms.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def world(group=None):
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def shard_rows(n_rows: int, rank: int, world_size: int):
    """Contiguous, balanced row block [lo, hi) of rank ``rank`` (first n % P ranks get one more row)."""
    base, extra = divmod(int(n_rows), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_replicates(job_indices, rank: int, world_size: int):
    """The reference's stride rule: job i goes to worker (i mod P) (cnmf.py:53-54)."""
    return [j for i, j in enumerate(job_indices) if (i - rank) % world_size == 0]


def allreduce_sum_(tensors, group=None):
    """In-place sum over ranks of a list of tensors (one flattened message per dtype, so the chunk's A and B
    travel together: R*Kp*(Kp+G)*4 bytes, 6 MB at K=15, R=50)."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return tensors
    by_dtype = {}
    for t in tensors:
        by_dtype.setdefault(t.dtype, []).append(t)
    for ts in by_dtype.values():
        flat = torch.cat([t.reshape(-1) for t in ts])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        off = 0
        for t in ts:
            n = t.numel()
            t.copy_(flat[off:off + n].view_as(t))
            off += n
    return tensors


def allreduce_flat_(t, group=None):
    """In-place sum over ranks of ONE contiguous tensor (the online loop's [A | B] message)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def allreduce_max_int(value: int, device, group=None) -> int:
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return int(value)
    t = torch.tensor([int(value)], dtype=torch.int64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return int(t.item())


def allreduce_scalars(values, device, group=None):
    """Sum of a few host doubles over ranks (||X||^2, sum X, row counts)."""
    v = np.asarray(values, dtype=np.float64)
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return v
    t = torch.from_numpy(v.copy()).to(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t.cpu().numpy()
