"""Probe: does running the LK batch as two half-batches on two CUDA streams overlap the latency-bound kernels of one half
with the ALU-bound FAST kernel of the other?  python tools/overlap_probe.py  (prints pairs/s for 1 x 512 and 2 x 256)."""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("aero-optical-flow_b200")
FX, FY, W, H, NF, ROT = 216.6677, 216.2457, 640, 480, 20, 4


def frames(n, seed):
    rng = np.random.default_rng(seed)
    seqs = [pkg.synth.lk_sequence(ROT, W, H, int(rng.integers(-3, 4)), int(rng.integers(-3, 4)), seed=seed + i) for i in range(8)]
    f = np.stack([np.stack([seqs[i % 8][k] for i in range(n)]) for k in range(ROT)])
    return torch.from_numpy(f).cuda()


def run(parts, n_each, steps=60, prio=None):
    flows, bufs, ts, outs, streams = [], [], [], [], []
    for p in range(parts):
        flows.append(pkg.LkFlow(FX, FY, -1, W, H, NF, n_streams=n_each))
        bufs.append(frames(n_each, 10 + p))
        ts.append(torch.zeros(n_each, dtype=torch.int32, device="cuda"))
        outs.append(torch.zeros((n_each, 4), dtype=torch.int32, device="cuda"))
        streams.append(torch.cuda.Stream(priority=(prio[p] if prio else 0)))
    torch.cuda.synchronize()

    def step(k):
        for p in range(parts):
            # triangle wave over the rotating frames so consecutive frames are consecutive in time
            idx = k % (2 * ROT - 2)
            idx = idx if idx < ROT else 2 * ROT - 2 - idx
            flows[p].calc_flow_device(bufs[p][idx].data_ptr(), ts[p].data_ptr(), outs[p].data_ptr(), stream=streams[p].cuda_stream)

    for k in range(6):
        step(k)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for p in range(parts):
        streams[p].wait_event(e0)
    for k in range(6, 6 + steps):
        step(k)
    for p in range(parts):
        e = torch.cuda.Event(); e.record(streams[p]); torch.cuda.current_stream().wait_event(e)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    for f in flows:
        f.close()
    return parts * n_each / (ms * 1e-3), ms


for parts, n_each, prio in ((1, 512, None), (2, 256, None), (4, 128, None), (2, 256, (0, -1))):
    v, ms = run(parts, n_each, prio=prio)
    print("%d x %d streams prio=%s: %.0f pairs/s, %.3f ms per 512" % (parts, n_each, prio, v, ms * 512 / (parts * n_each)))
