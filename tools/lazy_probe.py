#!/usr/bin/env python
"""Lazy-FAST statistics on the bench's LK workload: per step, bins picked lazily / second collecting passes / bins that
failed / stream-frames sent to the complete path (aof_lk_lazy_stats).   python tools/lazy_probe.py [streams] [steps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import bench
pkg = bench.entry.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
rot = bench.LK_ROT
frames, vel = bench.lk_host_frames(n, rot, seed=2000)
d = torch.from_numpy(frames.reshape(rot, -1)).cuda()
d_t = torch.zeros((rot, n), dtype=torch.int32, device="cuda")
d_out = torch.zeros((n, 4), dtype=torch.int32, device="cuda")
d_raw = torch.zeros((n, 6), dtype=torch.int32, device="cuda")
f = pkg.LkFlow(bench.FX, bench.FY, -1, bench.LK_W, bench.LK_H, bench.LK_FEATURES, n_streams=n)
last = (0, 0, 0, 0)
for k in range(steps):
    f.calc_flow_device(d[k % rot].data_ptr(), d_t[k % rot].data_ptr(), d_out.data_ptr(), d_raw.data_ptr())
    torch.cuda.synchronize()
    st = f.lazy_stats()
    raw = d_raw.cpu().numpy().view(pkg.capi.lk_raw_dtype).reshape(-1)
    print("step %d: lazy bins %d rebuilds %d failed bins %d streams->complete %d | bins needing FAST %.2f/16" %
          ((k,) + tuple(a - b for a, b in zip(st, last)) + (raw["fast_bins"].mean(),)))
    last = st
f.close()
