#!/usr/bin/env python
"""Pinned host -> device copy bandwidth with 1, 2, 4, 8 GPUs copying AT THE SAME TIME (the end-to-end
path is one H2D copy of the frames per step; its scaling is bounded by what the box's host side sustains).

One host thread per GPU.  Three placements of the pinned source buffers:
  main    allocated and first touched by the main thread (what a single-process caller gets by default)
  thread  allocated and first touched by the copying thread, no affinity
  local   the copying thread first pins itself to the CPUs the kernel lists as local to its GPU
          (/sys/bus/pci/devices/<bdf>/local_cpulist), then allocates and touches
Writes one JSON document (stdout, and the path given as argv[1]).  Run on the GPU box:  python tools/h2d_scaling.py out.json
"""
import json
import os
import subprocess
import sys
import threading
import time

import torch

MB = 256
REPS = 12


def local_cpus(dev):
    try:
        pr = torch.cuda.get_device_properties(dev)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
    except Exception:
        return None
    path = "/sys/bus/pci/devices/%s/local_cpulist" % bdf
    try:
        txt = open(path).read().strip()
    except OSError:
        return None
    cpus = set()
    for part in txt.split(","):
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        elif part:
            cpus.add(int(part))
    return sorted(cpus)


def run(n_gpus, placement, direction="h2d"):
    nbytes = MB << 20
    bufs = [None] * n_gpus
    if placement == "main":
        for i in range(n_gpus):
            bufs[i] = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
            bufs[i].fill_(i + 1)
    res = [None] * n_gpus
    start = threading.Barrier(n_gpus)

    def work(i):
        if placement == "local":
            cpus = local_cpus(i)
            if cpus:
                try:
                    os.sched_setaffinity(0, cpus)
                except OSError:
                    pass
        torch.cuda.set_device(i)
        if bufs[i] is None:
            bufs[i] = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
            bufs[i].fill_(i + 1)
        dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda:%d" % i)
        st = torch.cuda.Stream(device=i)
        src, dst = (bufs[i], dev) if direction == "h2d" else (dev, bufs[i])
        with torch.cuda.stream(st):
            for _ in range(2):
                dst.copy_(src, non_blocking=True)
            st.synchronize()
            start.wait()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(REPS):
                dst.copy_(src, non_blocking=True)
            e1.record(st)
            st.synchronize()
        res[i] = nbytes * REPS / (e0.elapsed_time(e1) * 1e-3) / 1e9

    th = [threading.Thread(target=work, args=(i,)) for i in range(n_gpus)]
    t0 = time.time()
    for t in th:
        t.start()
    for t in th:
        t.join()
    return {"gpus": n_gpus, "placement": placement, "direction": direction, "per_gpu_gbs": [round(r, 2) for r in res],
            "aggregate_gbs": round(sum(res), 2), "min_gbs": round(min(res), 2), "wall_s": round(time.time() - t0, 2)}


def sh(cmd):
    try:
        return subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=20).stdout
    except Exception as e:  # noqa: BLE001
        return "unavailable: %r" % (e,)


def main():
    n = torch.cuda.device_count()
    doc = {"devices": n, "copy_mb": MB, "reps": REPS, "host_cpus": os.cpu_count(),
           "affinity_of_this_process": sorted(os.sched_getaffinity(0)),
           "numa_nodes": sh("ls -d /sys/devices/system/node/node* 2>/dev/null | wc -l").strip(),
           "numa_cpulists": sh("cat /sys/devices/system/node/node*/cpulist 2>/dev/null").split(),
           "gpu_local_cpulist": {str(i): local_cpus(i) and "%d-%d (%d cpus)" % (min(local_cpus(i)), max(local_cpus(i)), len(local_cpus(i)))
                                 for i in range(n)},
           "topo": sh("nvidia-smi topo -m"), "lscpu": sh("lscpu | head -25"), "runs": []}
    counts = [c for c in (1, 2, 4, 8) if c <= n]
    for placement in ("main", "thread", "local"):
        for c in counts:
            doc["runs"].append(run(c, placement))
    for c in counts:
        doc["runs"].append(run(c, "local", "d2h"))
    txt = json.dumps(doc, indent=1)
    print(txt)
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(txt)


if __name__ == "__main__":
    main()
