"""N > 1 host logic on CPU: world_size-2 (and 3) gloo processes each own a contiguous range of streams, run
them independently (the oracle stands in for the GPU here -- there is none on this box), and the host gather
on rank 0 equals the single-process result, whatever the world size."""
import os
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FX, FY = 216.6677, 216.2457
TOTAL, T = 11, 4
out_dtype = np.dtype([("quality", np.int32), ("dt_us", np.int32), ("flow_x", np.float32), ("flow_y", np.float32)])


def run_streams(pkg, lo, hi):
    """calcFlow of streams [lo, hi) over T frames with per-stream oracle objects -> last frame's results"""
    from oracle import sad as osad
    port = osad.Port()
    res = np.zeros(hi - lo, out_dtype)
    for s in range(lo, hi):
        prev, cur, _ = pkg.synth.sad_pairs(T, max_shift=4, seed=1000 + s)
        o = port.px4(FX, FY, -1, 64, 64, 4, 30, 3000)
        for k in range(T):
            q, dt, fx, fy, *_ = o.calc_flow(cur[k] if k else prev[0], 1000 + 5000 * k)
        res[s - lo] = (q, dt, fx, fy)
    return res


def worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import __graft_entry__ as entry
    pkg = entry.load_package()
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = pkg.sharding.shard_range(TOTAL, world, rank)
    local = run_streams(pkg, lo, hi)
    dist.barrier()
    full = pkg.sharding.gather_results(local, TOTAL, world, rank, dist)
    # the non-blocking form (bench.py keeps one outstanding while the next step runs): two in flight, same answer
    h1 = pkg.sharding.gather_results_async(local, TOTAL, world, rank, dist)
    h2 = pkg.sharding.gather_results_async(local[::-1].copy()[::-1], TOTAL, world, rank, dist)
    a1, a2 = h1.wait(), h2.wait()
    if rank == 0:
        assert np.array_equal(a1, full) and np.array_equal(a2, full)
    else:
        assert a1 is None and a2 is None
    slowest = pkg.sharding.max_over_ranks(10.0 + rank, world, dist)
    if rank == 0:
        q.put((full, slowest))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_equals_single_process(pkg, world):
    want = run_streams(pkg, 0, TOTAL)
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=worker, args=(r, world, 29500 + world * 7 + os.getpid() % 500, q)) for r in range(world)]
    for p in procs:
        p.start()
    full, slowest = q.get()
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert np.array_equal(full, want)
    assert slowest == 10.0 + world - 1


def test_shard_ranges_partition(pkg):
    sh = pkg.sharding
    for total in (1, 7, 8, 4096, 65536 + 4096):
        for world in (1, 2, 3, 4, 8):
            edges = [sh.shard_range(total, world, r) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == total
            assert all(edges[r][1] == edges[r + 1][0] for r in range(world - 1))
            sizes = [b - a for a, b in edges]
            assert max(sizes) - min(sizes) <= 1
            for s in {0, total // 2, total - 1}:
                r = sh.owner_of(s, total, world)
                assert edges[r][0] <= s < edges[r][1]


def shm_worker(rank, world, path, q):
    sys.path.insert(0, ROOT)
    import time
    import __graft_entry__ as entry
    sh = entry.load_package().sharding
    if rank != 0:
        while not os.path.exists(path + ".ready"):
            time.sleep(0.01)
    g = sh.ShmGather(path, TOTAL, world, rank, out_dtype, create=rank == 0)
    if rank == 0:
        open(path + ".ready", "w").close()
    lo, hi = sh.shard_range(TOTAL, world, rank)
    ok = True
    for step in range(1, 40):
        local = np.zeros(hi - lo, out_dtype)
        local["quality"] = np.arange(lo, hi) + 1000 * step
        local["flow_x"] = rank + step
        g.put(local, step)
        if rank == 0 and step >= 2:                      # one step in flight, as bench.py does
            got = g.wait(step - 1)
            ok = ok and np.array_equal(got["quality"], np.arange(TOTAL) + 1000 * (step - 1))
            ok = ok and all(got["flow_x"][sh.shard_range(TOTAL, world, r)[0]] == r + step - 1 for r in range(world))
    if rank == 0:
        got = g.wait(39)
        ok = ok and np.array_equal(got["quality"], np.arange(TOTAL) + 39000)
        q.put(bool(ok))
    g.close()


@pytest.mark.parametrize("world", [2, 3])
def test_shared_memory_gather(pkg, world, tmp_path):
    path = str(tmp_path / "gather.bin")
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=shm_worker, args=(r, world, path, q)) for r in range(world)]
    for p in procs:
        p.start()
    assert q.get() is True
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
