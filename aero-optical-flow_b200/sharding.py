"""Whole-stream sharding across the GPUs of one box: every rank owns a contiguous, sticky
range of streams (their state lives on that GPU), there is no data-path collective, and the 16 B/stream
results are gathered on the host.  torch.distributed is plumbing only (barrier, max of the device time,
the gather of results); backend nccl on GPUs, gloo in the CPU tests."""
import numpy as np


def shard_range(total, world, rank):
    """contiguous [lo, hi) of `total` streams owned by `rank`; sizes differ by at most one"""
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def owner_of(stream, total, world):
    """rank that owns `stream` (inverse of shard_range)"""
    base, extra = divmod(total, world)
    cut = extra * (base + 1)
    return stream // (base + 1) if stream < cut else extra + (stream - cut) // max(base, 1)


def gather_results(local, total, world, rank, dist=None, group=None):
    """host gather: rank 0 receives every rank's structured result array into one array of `total` entries
    (None on the other ranks).  `local` = this rank's results for shard_range(total, world, rank)."""
    local = np.ascontiguousarray(local)
    lo, hi = shard_range(total, world, rank)
    assert len(local) == hi - lo
    if world == 1:
        return local
    parts = [None] * world if rank == 0 else None
    dist.gather_object(local, parts, dst=0, group=group)   # (bench.py: the gloo host group; the default group there is nccl)
    if rank != 0:
        return None
    out = np.empty(total, local.dtype)
    for r, p in enumerate(parts):
        a, b = shard_range(total, world, r)
        out[a:b] = p
    return out


def max_over_ranks(value, world, dist=None, device="cpu"):
    """max of a per-rank scalar (the benchmark's step time)"""
    if world == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
