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


class GatherHandle:
    """an outstanding host gather (gather_results_async); wait() -> the gathered array on rank 0, None elsewhere"""

    def __init__(self, work, parts, keep, total, world, rank, dtype):
        self.work, self.parts, self.keep, self.total, self.world, self.rank, self.dtype = work, parts, keep, total, world, rank, dtype

    def wait(self):
        if self.work is not None:
            self.work.wait()
        if self.rank != 0:
            return None
        if self.world == 1:
            return self.keep
        out = np.empty(self.total, self.dtype)
        for r, p in enumerate(self.parts):
            a, b = shard_range(self.total, self.world, r)
            out[a:b] = p.numpy().view(self.dtype)[: b - a]
        return out


def gather_results_async(local, total, world, rank, dist=None, group=None):
    """gather_results without blocking: the records travel as bytes (every rank pads its shard to the largest one) on
    the host group's own thread while the caller runs its next step; call .wait() on the returned handle."""
    import torch
    local = np.ascontiguousarray(local)
    lo, hi = shard_range(total, world, rank)
    assert len(local) == hi - lo
    if world == 1:
        return GatherHandle(None, None, local, total, world, rank, local.dtype)
    cap = -(-total // world)
    buf = np.zeros(cap, local.dtype)
    buf[: hi - lo] = local
    t = torch.from_numpy(buf.view(np.uint8).reshape(-1))
    parts = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
    work = dist.gather(t, parts, dst=0, group=group, async_op=True)
    return GatherHandle(work, parts, t, total, world, rank, local.dtype)


def max_over_ranks(value, world, dist=None, device="cpu"):
    """max of a per-rank scalar (the benchmark's step time)"""
    if world == 1:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


class ShmGather:
    """The host gather of one box without a collective: ONE array in POSIX shared memory (a file under /dev/shm) that
    every rank maps; a rank copies its records into its own slice and then publishes the step number in its own
    sequence slot; rank 0 copies the array out once every slot has reached the step and acknowledges it.  Two buffers
    alternate, so step k + 1 may be written while rank 0 still reads step k; put(k + 2) waits for the acknowledgement
    of step k.  x86 stores are not reordered with other stores, and numpy's copies are plain stores: the sequence
    number is written after the records.

    Every rank constructs it with the same `path`; rank 0 creates the file BEFORE the others open it (the caller
    places a barrier in between, bench.py does)."""

    def __init__(self, path, total, world, rank, dtype, create=False):
        self.total, self.world, self.rank, self.dtype = total, world, rank, np.dtype(dtype)
        self.rec_bytes = 2 * total * self.dtype.itemsize
        nbytes = 64 * world + self.rec_bytes                      # one cache line of sequence number per rank, then 2 buffers
        if create:
            with open(path, "wb") as f:
                f.truncate(nbytes)
        self.mm = np.memmap(path, dtype=np.uint8, mode="r+", shape=(nbytes,))
        self.seq = self.mm[: 64 * world].view(np.int64).reshape(world, 8)
        self.buf = self.mm[64 * world:].reshape(2, total * self.dtype.itemsize)   # bytes: numpy copies structured records field by field
        self.lo, self.hi = shard_range(total, world, rank)
        self.path = path

    def put(self, local, step, timeout_s=60.0):
        """this rank's records of `step` (1, 2, 3, ...).  Buffer step & 1 last held step - 2: the call waits until rank 0
        has acknowledged reading it."""
        assert len(local) == self.hi - self.lo
        import time
        t_end = time.monotonic() + timeout_s
        while self.rank != 0 and int(self.seq[0, 1]) < step - 2:
            if time.monotonic() > t_end:
                raise TimeoutError("ShmGather.put: step %d, rank 0 has read up to %d" % (step, int(self.seq[0, 1])))
        isz = self.dtype.itemsize
        self.buf[step & 1, self.lo * isz:self.hi * isz] = np.ascontiguousarray(local).view(np.uint8).reshape(-1)
        self.seq[self.rank, 0] = step

    def wait(self, step, timeout_s=60.0):
        """rank 0: a copy of the records of all ranks for `step` (and the acknowledgement that frees the buffer);
        other ranks: None"""
        if self.rank != 0:
            return None
        import time
        t_end = time.monotonic() + timeout_s
        while int(self.seq[:, 0].min()) < step:
            if time.monotonic() > t_end:
                raise TimeoutError("ShmGather: step %d, sequence numbers %s" % (step, self.seq[:, 0].tolist()))
        out = self.buf[step & 1].copy().view(self.dtype)
        self.seq[0, 1] = step
        return out

    def close(self, unlink=False):
        self.seq = self.buf = None
        del self.mm
        if unlink:
            import os
            try:
                os.unlink(self.path)
            except OSError:
                pass
