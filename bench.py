#!/usr/bin/env python
"""bench.py -- frame-pairs/s of the calcFlow hot path on N B200s (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload lk|sad|both] [--impl reference]

One process per GPU (torchrun for N>1).  A "step" is one calcFlow pass over one batch of
independent streams, each rank owning whole streams (weak scaling, no data-path collective;
torch.distributed is used only for the barrier and the max-over-ranks of the device time).

  lk  (primary, BASELINE configs[1] batched as in configs[4]): OpticalFlowOpenCV::calcFlow on
      640x480 frames, 20 features, 512 streams per GPU per launch.
  sad (BASELINE configs[2]): OpticalFlowPX4::calcFlow on 64x64 frames, search +-4, 4096 streams
      per GPU per launch.  With --workload both (default) it is reported inside the same JSON
      line under "sad".
Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402

FX, FY = 216.6677, 216.2457
# ---- SAD workload (config 5 = 65 536 streams over 8 GPUs = 8192 per GPU)
SAD_STREAMS, SAD_W, SAD_H, SAD_SEARCH, SAD_ROT = int(os.environ.get("AOF_BENCH_SAD_STREAMS", "4096")), 64, 64, 4, 16
SAD_STREAMS_CONFIG5 = 8192
SAD_BYTES_PER_PAIR = 2 * SAD_W * SAD_H + 16      # both frames read + 16 B result
SAD_INSTR_PER_PAIR = 40550                       # packed-byte instructions of the reference algorithm
# ---- LK workload (SURVEY.md 8(d), config 2 batched)
LK_STREAMS, LK_W, LK_H, LK_FEATURES, LK_ROT = 512, 640, 480, 20, 4
LK_LEVEL_PX = [640 * 480, 320 * 240, 160 * 120, 80 * 60]
LK_BYTES_PER_PAIR = LK_LEVEL_PX[0] + sum(LK_LEVEL_PX[1:])   # new frame read once + L1..L3 written = 408 000 B
# FAST response per pixel pair: 72 packed min/max (50 VIMNMX.U16x2 + 22 VIMNMX3.U16x2) on the raw ring values, all on the
# ALU pipe at 63 lanes/clk/SM (profiles/r1h_ubench_int_pipes.jsonl; the 125 of the earlier r1g file was ptxas fusing a chain
# of max into VIMNMX3)
LK_FAST_OPS_PER_PIXEL_PAIR = 72
MIN_TIMED_S = 0.5          # the K timed steps are repeated (each repeat bracketed like the first) until this much is timed


def workload_name(kind, n):
    """config.workload: ONE string per workload, the same in the b200 arm and in --impl reference"""
    if kind == "sad":
        return "sad64_search4_batch%d (BASELINE config 3: OpticalFlowPX4::calcFlow, %d 64x64 streams per GPU per step)" % (n, n)
    return ("lk640x480_features20_batch%d (BASELINE config 2 batched as in config 5: OpticalFlowOpenCV::calcFlow, "
            "%d 640x480 streams per GPU per step, FAST + binned selection + 3-level pyramidal LK 9x9)" % (n, n))


def ncu_traffic(kernel, n, n_ref):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (profiles/traffic.json), scaled from the
    capture's streams per launch (n_ref) to this run's; None when there is no capture."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return float(t[kernel]) * n / n_ref
    except Exception:
        return None


def peaks():
    p = {"hbm_gbs": 6650.0, "which": "fallback"}
    try:
        m = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        p = {"hbm_gbs": float(m["hbm_gbs"]), "which": "measured"}
    except Exception:
        pass
    try:
        a = json.load(open(os.path.join(ROOT, "profiles", "alu_peak.json")))
        p["alu_packed_instr_per_s"] = float(a["vabsdiff4_thread_ops_per_s"])
        p["alu_which"] = a.get("which", "measured")
    except Exception:
        pass
    return p


class ClockSampler:
    """nvidia-smi clocks / throttle reasons, sampled every 20 ms with nvidia-smi's own timestamps; stop(t0, t1) keeps the
    samples that fall inside [t0, t1] (the timed region, time.time() seconds)."""
    Q = ("timestamp,clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self, wait_s=4.0):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, bufsize=1)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
            t_end = time.time() + wait_s                      # nvidia-smi needs a moment before its first sample
            while not self.rows and time.time() < t_end:
                time.sleep(0.02)
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    @staticmethod
    def _stamp(text, fallback):
        import datetime
        try:
            return datetime.datetime.strptime(text.strip(), "%Y/%m/%d %H:%M:%S.%f").timestamp()
        except Exception:
            return fallback

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, sm_all = [], [], set(), []
        for t_read, r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                clk, cmax = float(f[1]), float(f[2])
            except ValueError:
                continue
            sm_all.append(clk)
            t = self._stamp(f[0], t_read)
            if t0 is not None and not (t0 - 0.01 <= t <= t1 + 0.01):
                continue
            sm.append(clk); mx.append(cmax)
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "samples_outside_timed_region": len(sm_all) - len(sm),
                "window": "samples whose nvidia-smi timestamp lies inside the timed region"}


def spread_device(local_rank, n_devices):
    """local rank -> device, alternating between the two halves of the box (0, n/2, 1, n/2 + 1, ...).  On the 8-GPU boxes of
    this pool the pinned host -> device path is shared per group of four GPUs (profiles/r2_h2d_scaling.json: 52.6 GB/s for
    one GPU, 2 x 55 for GPUs 0-1, but 4 x 29 for GPUs 0-3; at 8 GPUs 4 x 23.5 + 4 x 35.8): ranks 0..N-1 on devices 0..N-1
    put all of a 4-rank job behind one group.  Identity when N ranks see N devices."""
    if n_devices < 4 or n_devices % 2 or local_rank >= n_devices:
        return local_rank
    half = n_devices // 2
    return (local_rank // 2) + (local_rank % 2) * half


def dist_setup():
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = spread_device(int(os.environ.get("LOCAL_RANK", "0")), torch.cuda.device_count())
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if not dist.is_initialized():
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        global HOST_GROUP
        HOST_GROUP = dist.new_group(backend="gloo")          # the host-side gather of the results (no data-path collective)
    return rank, world, local


HOST_GROUP = None


_GATHER_SEQ = [0]


def host_gather_open(sample, rank, world):
    """The 'host gather' of SURVEY.md 8(e) for the ranks of one box: one array in POSIX shared memory that every rank
    copies its 16-byte records into and rank 0 reads (aero-optical-flow_b200/sharding.py: ShmGather, covered on CPU by
    tests/test_sharding.py) -- host memory only, no collective.  (The gloo form, sharding.gather_results, costs a
    collective's latency per step, more than a whole 0.38 ms SAD step.)  -> gather object, None for one rank"""
    if world == 1:
        return None
    import torch.distributed as dist
    sh = entry.load_package().sharding
    _GATHER_SEQ[0] += 1
    path = "/dev/shm/aof_bench_gather_%s_%d.bin" % (os.environ.get("MASTER_PORT", "0"), _GATHER_SEQ[0])
    g = sh.ShmGather(path, world * len(sample), world, rank, sample.dtype, create=True) if rank == 0 else None
    dist.barrier(group=HOST_GROUP)                               # the file exists before the others map it
    if rank != 0:
        g = sh.ShmGather(path, world * len(sample), world, rank, sample.dtype)
    return g


def barrier(world):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(ms, world):
    import torch
    if world == 1:
        return ms
    import torch.distributed as dist
    t = torch.tensor([ms], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def time_steps(step, first, steps, world, min_s=MIN_TIMED_S, max_repeats=400):
    """K = `steps` calls of step(k) between a barrier + torch.cuda.synchronize() on both sides, device time from CUDA
    events on the launching stream, max over ranks.  The K-step region is repeated, every repeat bracketed in the same
    way, until at least min_s of device time has been timed (a 10 ms region is too short for a clock sampler).
    -> {"ms": mean ms per K steps, "repeats": R, "all_ms": [...], "t0", "t1": wall-clock bounds, "next": next k}"""
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    all_ms, k, repeats = [], first, 1
    t0 = time.time()
    while len(all_ms) < repeats:
        barrier(world)
        e0.record()
        for _ in range(steps):
            step(k); k += 1
        e1.record()
        torch.cuda.synchronize()
        all_ms.append(max_over_ranks(e0.elapsed_time(e1), world))       # the same number on every rank
        if len(all_ms) == 1:
            repeats = int(min(max_repeats, max(1, np.ceil(min_s * 1e3 / max(all_ms[0], 1e-3)))))
    barrier(world)
    return {"ms": float(np.mean(all_ms)), "repeats": repeats, "all_ms": all_ms, "t0": t0, "t1": time.time(), "next": k}


# ----------------------------------------------------------------------------- synthetic inputs
def sad_host_frames(n, rot, seed):
    """rot consecutive frames of n translating 64x64 streams (numpy, host)."""
    pkg = entry.load_package()
    rng = np.random.default_rng(seed)
    span = SAD_SEARCH * (rot // 2) + 8
    tex = pkg.synth.box_blur3(rng.integers(0, 256, (n, SAD_H + 2 * span, SAD_W + 2 * span), dtype=np.uint8))
    vel = rng.integers(-SAD_SEARCH, SAD_SEARCH + 1, (n, 2))
    frames = np.empty((rot, n, SAD_H, SAD_W), np.uint8)
    for k in range(rot):
        for i in range(n):
            p = k if k <= rot // 2 else rot - k      # triangle wave: every consecutive pair (and the wrap) moves by +-vel
            ox, oy = span - p * vel[i, 0], span - p * vel[i, 1]
            frames[k, i] = tex[i, oy:oy + SAD_H, ox:ox + SAD_W]
    return frames, vel


def lk_host_frames(n, rot, seed, out=None):
    """rot frames of n translating 640x480 streams: 8 Gaussian-blurred noise textures (sigma 2, as
    SURVEY.md 8(d)), each stream a different window and velocity; triangle-wave positions so that
    every consecutive pair, and the wrap-around, is a pure translation by +-vel."""
    from scipy.ndimage import gaussian_filter
    rng = np.random.default_rng(seed)
    n_tex, m = 8, 64
    tex = []
    for t in range(n_tex):
        a = gaussian_filter(rng.random((LK_H + 2 * m + 64, LK_W + 2 * m + 64)), 2.0, mode="wrap")
        a -= a.min(); a /= a.max()
        tex.append(np.round(a * 255.0).astype(np.uint8))
    vel = rng.integers(-5, 6, (n, 2))
    org = rng.integers(0, 64, (n, 2))
    frames = out if out is not None else np.empty((rot, n, LK_H, LK_W), np.uint8)
    for k in range(rot):
        p = k if k <= rot // 2 else rot - k
        for i in range(n):
            ox, oy = m + org[i, 0] - p * vel[i, 0], m + org[i, 1] - p * vel[i, 1]
            frames[k, i] = tex[i % n_tex][oy:oy + LK_H, ox:ox + LK_W]
    return frames, vel


# ----------------------------------------------------------------------------- CPU arms
def cpu_sad_runner():
    """-> (fn(prev, cur, threads), kind): the unmodified reference objects when oracle/_ref was built"""
    from oracle import sad as osad
    if osad.ref_available():
        R = osad.Ref()
        return (lambda p, c, th: R.compute_flow_batch(p, c, SAD_SEARCH, 30, 3000, threads=th)), "reference"
    P = osad.Port()
    return (lambda p, c, th: P.compute_flow_batch(p, c, SAD_SEARCH, 30, 3000)), "port"


def cpu_sad(frames, budget_s):
    fn, kind = cpu_sad_runner()
    cores = os.cpu_count() or 1
    threads = cores if kind == "reference" else 1
    n = frames.shape[1]
    t0 = time.perf_counter(); fn(frames[0][:256], frames[1][:256], threads); probe = time.perf_counter() - t0
    per_pair = max(probe / 256, 1e-7)
    reps = int(max(1, min(64, budget_s / (per_pair * n))))
    t0 = time.perf_counter()
    for r in range(reps):
        fn(frames[r % (len(frames) - 1)], frames[r % (len(frames) - 1) + 1], threads)
    dt = time.perf_counter() - t0
    return {"value": reps * n / dt, "unit": "frame-pairs/s", "cores": threads, "kind": kind,
            "sample": "%d x %d SAD 64x64 +-%d pairs, PX4Flow::compute_flow (%s), %d host threads" % (reps, n, SAD_SEARCH, kind, threads)}


def _lk_cpu_tracker_factory():
    """-> (make(), kind, what): the reference's own OpticalFlowOpenCV object where it can run --
    oracle/_ref/libaof_ref_lk.so = trackFeatures.cpp + flow_opencv.cpp + optical_flow.cpp compiled unmodified, calling
    the real OpenCV (cv2: FAST, calcOpticalFlowPyrLK) the way the reference links it -- else the same objects over
    oracle/lk_oracle.c's restatement of those three OpenCV calls, else the C port alone."""
    from oracle import ref_lk
    if ref_lk.available():
        try:
            import cv2  # noqa: F401
            return (lambda: ref_lk.RefTracker(FX, FY, -1, LK_W, LK_H, LK_FEATURES, backend="cv2")), "reference", \
                "the reference's trackFeatures.cpp + flow_opencv.cpp objects (oracle/_ref) over OpenCV %s via cv2, 1 OpenCV thread per process" % cv2.__version__
        except Exception:
            return (lambda: ref_lk.RefTracker(FX, FY, -1, LK_W, LK_H, LK_FEATURES)), "port", \
                "the reference's trackFeatures.cpp + flow_opencv.cpp objects (oracle/_ref) over oracle/lk_oracle.c's FAST / PyrLK"
    from oracle import lk as olk
    return (lambda: olk.Tracker(FX, FY, -1, LK_W, LK_H, LK_FEATURES)), "port", "oracle/lk_oracle.c (CPU restatement of the whole path)"


def _lk_cpu_worker(wid, procs, path, shape, conn):
    """one process = the trackers of streams wid, wid + procs, ...; a step = calcFlow of ONE frame on each of them"""
    try:
        devnull = os.open(os.devnull, os.O_WRONLY)           # the reference prints from calcFlow: keep bench's stdout to its ONE line
        os.dup2(devnull, 1)
        make, _, _ = _lk_cpu_tracker_factory()
        frames = np.memmap(path, dtype=np.uint8, mode="r", shape=shape)
        rot, n = frames.shape[:2]
        mine = list(range(wid, n, procs))
        trackers = [make() for _ in mine]
        for t, i in zip(trackers, mine):
            t.calc_flow(frames[0, i], 0)                     # priming frame (detection only), untimed
        conn.send(("ready", len(mine)))
        while True:
            k = conn.recv()
            if k is None:
                break
            q_ok = 0
            for t, i in zip(trackers, mine):
                q = t.calc_flow(frames[k % rot, i], 1000 * k)[0]
                q_ok += q > 60
            conn.send(("done", q_ok))
    except Exception as e:                                   # noqa: BLE001
        conn.send(("error", repr(e)))


class CpuLk:
    """The LK workload on the host cores: one process per core (the reference's tracker keeps its state in file statics,
    and the cv2 callbacks hold the GIL, so threads would serialise), whole streams per process, a step = one calcFlow
    pass over all n streams -- the same step as the GPU arm's."""

    def __init__(self, frames, procs=None):
        import multiprocessing as mp
        import tempfile
        self.n = frames.shape[1]
        self.procs = min(procs or (os.cpu_count() or 1), self.n)
        _, self.kind, self.what = _lk_cpu_tracker_factory()
        # the frames go through a file in shared memory: workers are SPAWNED (a fork of a process that holds a CUDA
        # context and its threads is not safe), and page-locked memory does not survive a fork anyway
        shm = "/dev/shm" if os.path.isdir("/dev/shm") else None
        fd, self.path = tempfile.mkstemp(prefix="aof_bench_lk_", suffix=".u8", dir=shm)
        os.close(fd)
        mm = np.memmap(self.path, dtype=np.uint8, mode="w+", shape=frames.shape)
        mm[:] = frames
        mm.flush()
        del mm
        ctx = mp.get_context("spawn")
        self.workers = []
        for w in range(self.procs):
            a, b = ctx.Pipe()
            pr = ctx.Process(target=_lk_cpu_worker, args=(w, self.procs, self.path, tuple(frames.shape), b), daemon=True)
            pr.start()
            self.workers.append((pr, a))
        for _, c in self.workers:
            tag, v = c.recv()
            if tag != "ready":
                raise RuntimeError("cpu LK worker: %s" % (v,))
        self.k = 1

    def step(self):
        """-> seconds of wall clock for one pass over the n streams"""
        t0 = time.perf_counter()
        for _, c in self.workers:
            c.send(self.k)
        ok = 0
        for _, c in self.workers:
            tag, v = c.recv()
            if tag != "done":
                raise RuntimeError("cpu LK worker: %s" % (v,))
            ok += v
        dt = time.perf_counter() - t0
        self.k += 1
        assert ok > 0.9 * self.n, "CPU LK arm lost track (%d of %d streams with quality > 60)" % (ok, self.n)
        return dt

    def close(self):
        for pr, c in self.workers:
            try:
                c.send(None)
            except Exception:
                pass
        for pr, _ in self.workers:
            pr.join(timeout=5)
            if pr.is_alive():
                pr.terminate()
        try:
            os.unlink(self.path)
        except OSError:
            pass


def cpu_lk(frames, budget_s):
    """cpu_baseline of the LK workload: a bounded sample (whole passes over the batch until budget_s is spent, at least 2)"""
    arm = CpuLk(frames)
    try:
        times = [arm.step()]
        while sum(times) < budget_s and len(times) < 50 or len(times) < 2:
            times.append(arm.step())
    finally:
        arm.close()
    med = float(np.median(times))
    return {"value": arm.n / med, "unit": "frame-pairs/s", "cores": arm.procs, "kind": arm.kind,
            "sample": "%d passes over %d streams (640x480, 20 features), median pass %.3f s; %s; %d processes" %
                      (len(times), arm.n, med, arm.what, arm.procs)}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores (rank 0 only), on the b200 arm's
    own workload: a step = one calcFlow pass over the same batch of streams."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    entry.build_oracles()
    total = max(1, args.steps + args.warmup)
    if args.workload == "sad":
        n = SAD_STREAMS
        frames, _ = sad_host_frames(n, 3, 1)
        fn, kind = cpu_sad_runner()
        cores = os.cpu_count() or 1
        threads = cores if kind == "reference" else 1
        times = []
        for k in range(total):
            t0 = time.perf_counter(); fn(frames[k % 2], frames[k % 2 + 1], threads); times.append(time.perf_counter() - t0)
        times = times[args.warmup:]
        what = "PX4Flow::compute_flow of the reference (oracle/_ref), %d host threads" % threads if kind == "reference" else \
            "oracle/sad_oracle.c (CPU restatement), 1 thread"
        workload = workload_name("sad", n)
    else:
        n = args.lk_streams
        frames, _ = lk_host_frames(n, LK_ROT, 2000)
        arm = CpuLk(frames)
        try:
            times = [arm.step() for _ in range(total)][args.warmup:]
        finally:
            arm.close()
        kind, threads, what = arm.kind, arm.procs, arm.what
        workload = workload_name("lk", n)
    ms = 1e3 * float(np.mean(times))
    v = n / (ms * 1e-3)
    cb = {"value": v, "unit": "frame-pairs/s", "cores": threads, "kind": kind,
          "sample": "%d steps of %d streams each (%d stream-frames), mean step %.1f ms; %s" % (len(times), n, len(times) * n, ms, what)}
    print(json.dumps({
        "impl": "reference", "metric": "frame-pairs/s", "value": v, "unit": "frame-pairs/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload, "streams_per_step": n,
                   "note": "host cores only (the reference has no GPU path): one step = one calcFlow pass over the batch; "
                           "the value does not grow with --gpus"},
        "cpu_baseline": cb, "e2e": {"value": v, "unit": "frame-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ----------------------------------------------------------------------------- GPU arms
def kernel_breakdown(pkg, step, first, steps):
    """per-kernel device time of `steps` steps through the library's own event timing"""
    import torch
    L = pkg.lib()
    L.aof_timing_reset(); L.aof_timing_enable(1)
    for k in range(steps):
        step(first + k)
    torch.cuda.synchronize()
    times = pkg.capi.kernel_times()
    L.aof_timing_enable(0)
    return {name: {"ms_per_step": ms / steps, "launches_per_step": cnt / steps} for name, (ms, cnt) in times.items()}


def pinned_frames(shape):
    import torch
    t = torch.empty(shape, dtype=torch.uint8, pin_memory=True)
    return t, t.numpy()


def h2d_probe(keep, world):
    """Pinned host -> device bandwidth of this box as the end-to-end path sees it: one cudaMemcpyAsync of up to 256 MiB of the
    bench's own pinned frames, all ranks at the same time (they share the host side), best of 5, GB/s of THIS rank -> the
    min over ranks (the slowest link bounds the max-over-ranks e2e time)."""
    import torch
    src = keep.reshape(-1)[: 256 << 20]
    dst = torch.empty_like(src, device="cuda")
    best = 0.0
    for _ in range(5):
        barrier(world)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); dst.copy_(src, non_blocking=True); e1.record()
        torch.cuda.synchronize()
        best = max(best, src.numel() / (e0.elapsed_time(e1) * 1e-3) / 1e9)
    del dst
    return -max_over_ranks(-best, world)


def time_e2e(call, first, steps, world, rank, min_s=MIN_TIMED_S):
    """host-API steps (synchronous calls: H2D, kernels, D2H inside) between barriers, wall clock, max over ranks; every
    step's records are gathered on rank 0 through host shared memory (host_gather_open), rank 0 reading step k - 1 after
    its own step k, the last step awaited inside the timed region.  The step count is raised until min_s is timed.
    -> (ms total, steps timed, last gathered result on rank 0)"""
    import torch
    import torch.distributed as dist
    barrier(world)
    t0 = time.perf_counter(); res = call(first); probe = time.perf_counter() - t0
    probe = max_over_ranks(probe * 1e3, world) * 1e-3
    steps = int(max(steps, min(2000, np.ceil(min_s / max(probe, 1e-6)))))
    g = host_gather_open(res, rank, world)
    barrier(world)
    t0 = time.perf_counter()
    for k in range(steps):
        res = call(first + 1 + k)
        if g is not None:
            g.put(res, k + 1)
            if k >= 1:
                g.wait(k)
    if g is not None:
        out = g.wait(steps)
        res = out if rank == 0 else None
    torch.cuda.synchronize()
    ms = max_over_ranks((time.perf_counter() - t0) * 1e3, world)
    if g is not None:
        dist.barrier(group=HOST_GROUP)
        g.close(unlink=rank == 0)
    return ms, steps, res


def host_threaded_run(make, frames, n, n_devices, min_s=MIN_TIMED_S):
    """n_devices x n streams end to end through aof_*_create_multi from THIS process: pinned host frames in, every device's
    records in the library's one pinned array out.  -> dict"""
    import torch
    devices = [spread_device(r, torch.cuda.device_count()) for r in range(n_devices)]
    rot = min(2, frames.shape[0])
    per = frames[0].size // n                                            # bytes per frame
    keep = torch.empty((rot, n_devices * n, per), dtype=torch.uint8, pin_memory=True)
    host = keep.numpy()
    for k in range(rot):
        for d in range(n_devices):
            host[k, d * n:(d + 1) * n] = frames[k].reshape(n, per)
    m = make(n_devices * n, devices)
    t_us = np.zeros(n_devices * n, np.uint32)
    for k in range(3):
        m.calc_flow_batch(host[k % rot], t_us + 1000 * k, copy=False)
    t0 = time.perf_counter(); m.calc_flow_batch(host[1 % rot], t_us + 3000, copy=False); probe = time.perf_counter() - t0
    steps = int(max(3, min(2000, np.ceil(min_s / max(probe, 1e-6)))))
    t0 = time.perf_counter()
    for k in range(steps):
        res = m.calc_flow_batch(host[k % rot], t_us + 1000 * (k + 4), copy=False)
    dt = time.perf_counter() - t0
    assert res.size == n_devices * n and (res["quality"] > 60).mean() > 0.97
    out = {"value": n_devices * n * steps / dt, "unit": "frame-pairs/s", "steps": steps, "ms_per_step": dt / steps * 1e3,
           "devices": devices, "ranges": m.ranges(), "h2d_bytes_per_step": n_devices * n * per + 4 * n_devices * n,
           "d2h_bytes_per_step": 16 * n_devices * n,
           "api": "aof_*_create_multi + aof_*_multi_calc_flow_batch: one process, a host worker thread per device, results in the "
                  "library's one pinned array (host gather inside the call)"}
    m.close()
    del keep
    return out


def host_threaded_leg(make, frames, n, world, rank, min_s=MIN_TIMED_S):
    """The same end-to-end workload through the library's own multi-GPU entry points (aof_*_create_multi: ONE process, one
    host worker thread per device, every device's D2H landing in one pinned result array = the host gather, no torch and no
    collective).  Rank 0 drives all `world` devices while the other ranks wait at a barrier.  -> dict or None"""
    import torch
    if world == 1:
        return None
    import torch.distributed as dist
    torch.cuda.synchronize()
    dist.barrier(group=HOST_GROUP)          # host-side: the waiting ranks must not spin in a kernel on the GPUs rank 0 is about to use
    out = host_threaded_run(make, frames, n, world, min_s) if rank == 0 else None
    dist.barrier(group=HOST_GROUP)
    return out


def run_single_process(args):
    """python bench.py --gpus N without torchrun: ONE process drives N GPUs through aof_*_create_multi (host worker threads
    inside the library, no torch.distributed, no NCCL).  Only the end-to-end metric exists in this mode (the entry points
    take host frames); the device-resident value and the rooflines come from the per-GPU arm (torchrun or --gpus 1)."""
    pkg = entry.load_package()
    res = {}
    if args.workload in ("lk", "both"):
        n = args.lk_streams
        keep, h_frames = pinned_frames((2, n, LK_H, LK_W))
        lk_host_frames(n, 2, seed=2000, out=h_frames)
        res["lk"] = host_threaded_run(lambda total, devs: pkg.LkFlowMulti(FX, FY, -1, LK_W, LK_H, LK_FEATURES, n_streams=total, devices=devs),
                                      h_frames, n, args.gpus)
        res["lk"]["workload"] = workload_name("lk", n)
    if args.workload in ("sad", "both"):
        n = SAD_STREAMS
        frames, _ = sad_host_frames(n, 2, seed=1000)
        res["sad"] = host_threaded_run(lambda total, devs: pkg.Px4FlowMulti(FX, FY, -1, SAD_W, SAD_H, SAD_SEARCH, 30, 3000, n_streams=total, devices=devs),
                                       frames, n, args.gpus)
        res["sad"]["workload"] = workload_name("sad", n)
    p = res["sad" if args.workload == "sad" else "lk"]
    line = {"metric": "frame-pairs/s", "value": p["value"], "unit": "frame-pairs/s", "n_gpus": args.gpus, "steps": p["steps"],
            "warmup": 4, "ms_per_step": p["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": p["workload"], "mode": "single process, host-threaded multi-GPU entry points (aof_*_create_multi); "
                       "value == e2e: host frames in, gathered records out", "devices": p["devices"]},
            "roofline": None,
            "e2e": {"value": p["value"], "unit": "frame-pairs/s", "h2d_bytes_per_step": p["h2d_bytes_per_step"],
                    "d2h_bytes_per_step": p["d2h_bytes_per_step"], "api": p["api"]},
            "gpu_launches": int(pkg.lib().aof_launch_count())}
    if args.workload == "both":
        line["sad"] = res["sad"]
    print(json.dumps(line))


def find_kernel(kb, prefix):
    """the entry of the timing table whose kernel symbol starts with `prefix` (the table names what actually ran)"""
    for name in kb:
        if name.startswith(prefix):
            return name
    raise KeyError("no kernel %s* in %s" % (prefix, sorted(kb)))


def run_sad_gpu(args, rank, world, local, with_cpu, n_streams=None, light=False):
    """light: the throughput legs only (timed region + end to end), for the extra config-5 size"""
    import torch
    pkg = entry.load_package()
    L = pkg.lib()
    n, rot = n_streams or SAD_STREAMS, SAD_ROT
    frames, vel = sad_host_frames(n, rot, seed=1000 + rank)
    keep, h_frames = pinned_frames(frames.shape)
    h_frames[:] = frames
    fb = n * SAD_W * SAD_H
    d_frames = torch.empty((rot, fb), dtype=torch.uint8, device="cuda")
    d_frames.copy_(keep.reshape(rot, fb))
    d_t = torch.zeros((rot, n), dtype=torch.int32, device="cuda")
    d_out = torch.zeros((n, 4), dtype=torch.int32, device="cuda")
    flow = pkg.Px4Flow(FX, FY, -1, SAD_W, SAD_H, SAD_SEARCH, 30, 3000, n_streams=n, device=local)
    stream = torch.cuda.current_stream().cuda_stream
    fptr = [d_frames[k].data_ptr() for k in range(rot)]
    tptr = [d_t[k].data_ptr() for k in range(rot)]

    def step(k):
        flow.calc_flow_device(fptr[k % rot], tptr[k % rot], d_out.data_ptr(), stream=stream)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for k in range(args.warmup + 1):          # +1: the priming frame launches nothing
        step(k)
    l0 = L.aof_launch_count()
    tm = time_steps(step, args.warmup + 1, args.steps, world)
    launches = (L.aof_launch_count() - l0) // tm["repeats"]
    ms, k = tm["ms"], tm["next"]
    clocks = sampler.stop(tm["t0"], tm["t1"]) if rank == 0 else None
    out = d_out.cpu().numpy().view(pkg.capi.flow_out_dtype).reshape(-1)
    assert (out["quality"] == 255).mean() > 0.99, "timed SAD path produced implausible quality"
    kb = kernel_breakdown(pkg, step, k, min(args.steps, 200))

    # end to end through the host API: pinned host frames -> H2D -> kernels -> 16 B/stream D2H
    flow_e = pkg.Px4Flow(FX, FY, -1, SAD_W, SAD_H, SAD_SEARCH, 30, 3000, n_streams=n, device=local)
    t_us = np.zeros(n, np.uint32)
    for k in range(4):
        flow_e.calc_flow_batch(h_frames[k % rot], t_us + 1000 * k)
    e2e_ms, e_steps, res = time_e2e(lambda j: flow_e.calc_flow_batch(h_frames[(j + 4) % rot], t_us + 1000 * (j + 4)),
                                    0, min(args.steps, 200), world, rank)
    if rank == 0:
        assert res.size == world * n and (res["quality"] == 255).mean() > 0.99
    flow_e.close(); flow.close()
    h2d_gbs = h2d_probe(keep, world)
    multi = None if light else host_threaded_leg(
        lambda total, devs: pkg.Px4FlowMulti(FX, FY, -1, SAD_W, SAD_H, SAD_SEARCH, 30, 3000, n_streams=total, devices=devs),
        h_frames, n, world, rank)

    # BASELINE config 1 as written: ONE 64x64 stream, host frame in, result out (latency, not throughput)
    lat_us = None
    if not light:
        one = pkg.Px4Flow(FX, FY, -1, SAD_W, SAD_H, SAD_SEARCH, 30, 3000, n_streams=1, device=local)
        for k in range(4):
            one.calc_flow(h_frames[k % rot, 0], 1000 * k)
        t0 = time.perf_counter()
        for k in range(100):
            one.calc_flow(h_frames[(k + 4) % rot, 0], 1000 * (k + 4))
        lat_us = (time.perf_counter() - t0) / 100 * 1e6
        one.close()

    # BASELINE config 4: dense block field of 1280x720 pairs, +-8 (device-resident pairs, 8 distinct pairs rotating)
    dense = None
    if rank == 0 and not light:
        dn, dw, dh, ds = 8, 1280, 720, 8
        rng = np.random.default_rng(77)
        tex = pkg.synth.box_blur3(rng.integers(0, 256, (dh + 32, dw + 32), dtype=np.uint8))
        shifts = rng.integers(-ds, ds + 1, (dn, 2))
        d_prev = torch.from_numpy(np.stack([tex[16:16 + dh, 16:16 + dw]] * dn)).cuda()
        d_cur = torch.from_numpy(np.stack([tex[16 - sy:16 - sy + dh, 16 - sx:16 - sx + dw] for sx, sy in shifts])).cuda()
        df = pkg.SadDense(dw, dh, ds, 30, 3000, device=local)
        d_field = torch.zeros((dn, df.nby, df.nbx, pkg.capi.dense_block_dtype.itemsize), dtype=torch.uint8, device="cuda")
        run = lambda: df.compute_device(d_prev.data_ptr(), d_cur.data_ptr(), dn, d_field.data_ptr(), stream=stream)
        for _ in range(3):
            run()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        ev0.record()
        reps = 20
        for _ in range(reps):
            run()
        ev1.record()
        torch.cuda.synchronize()
        d_ms = ev0.elapsed_time(ev1) / reps
        field = d_field.cpu().numpy().view(pkg.capi.dense_block_dtype).reshape(dn, df.nby, df.nbx)
        hit = float(np.mean([((field[i]["dx"] == shifts[i, 0]) & (field[i]["dy"] == shifts[i, 1]) & (field[i]["flags"] == 3)).mean()
                             for i in range(dn)]))
        assert hit > 0.9, "dense SAD field does not recover the translations (%.3f)" % hit
        blocks = df.nbx * df.nby
        dense = {"workload": "dense_sad_1280x720_search8 (BASELINE config 4), %d pairs per launch" % dn, "pairs_per_s": dn / (d_ms * 1e-3),
                 "ms_per_pair": d_ms / dn, "blocks_per_pair": blocks, "translation_recovered_frac": float(hit),
                 "l2": "8 pairs = 14.7 MB: L2-resident, the kernel is integer-ALU bound by 6x (SURVEY 8d)"}
        df.close()
    if rank != 0:
        return None
    pk = peaks()
    dom = max(kb, key=lambda kname: kb[kname]["ms_per_step"])
    kname = find_kernel(kb, "sad_flow")                       # the SAD kernel that ran (sad_flow3_kernel at 64x64 +-4)
    kern_ms = kb[kname]["ms_per_step"] / max(1.0, kb[kname]["launches_per_step"])
    hbm = n * SAD_BYTES_PER_PAIR / (kern_ms * 1e-3) / 1e9
    hbm_obj = {"bound": "hbm", "achieved": hbm, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": hbm / pk["hbm_gbs"],
               "peak_source": pk["which"], "algorithmic_bytes_per_launch": n * SAD_BYTES_PER_PAIR,
               "traffic": ncu_traffic(kname.split("<")[0], n, 4096)}
    if "alu_packed_instr_per_s" in pk:
        # the binding roofline: 40 550 packed-byte SAD instructions per pair against the measured
        # VABSDIFF4 issue rate of the ALU pipe (profiles/alu_peak.json, tools/ubench.cu)
        a = n * SAD_INSTR_PER_PAIR / (kern_ms * 1e-3)
        roof = {"bound": "int-alu", "achieved": a, "peak": pk["alu_packed_instr_per_s"], "unit": "packed-byte instr/s",
                "frac": a / pk["alu_packed_instr_per_s"], "peak_source": pk["alu_which"] + " (tools/ubench.cu, VABSDIFF4 issue rate)",
                "algorithmic_instr_per_pair": SAD_INSTR_PER_PAIR, "algorithmic_instr_per_launch": n * SAD_INSTR_PER_PAIR,
                "traffic": hbm_obj["traffic"], "hbm": hbm_obj}
    else:
        roof = dict(hbm_obj, note="profiles/alu_peak.json missing: the binding integer-ALU roofline cannot be stated")
    roof.update({"kernel": kname, "kernel_ms": kern_ms, "dominant_kernel_by_time": dom})
    e2e_v = world * n * e_steps / (e2e_ms * 1e-3)
    e2e_gbs = (fb + 4 * n) * e_steps / (e2e_ms * 1e-3) / 1e9
    r = {
        "value": world * n * args.steps / (ms * 1e-3), "ms_per_step": ms / args.steps,
        "timed_region": {"repeats": tm["repeats"], "steps_per_repeat": args.steps, "total_s": sum(tm["all_ms"]) * 1e-3,
                         "ms_per_step_min": min(tm["all_ms"]) / args.steps, "ms_per_step_max": max(tm["all_ms"]) / args.steps},
        "config": {"workload": workload_name("sad", n),
                   "streams_per_gpu": n, "image": "64x64", "search": SAD_SEARCH, "output_rate": -1,
                   "l2": "current frames rotate over %d buffers = %d MiB > 126 MB L2; the previous frame is last step's "
                         "current frame by construction of calcFlow" % (rot, rot * fb >> 20),
                   "parallelism": "streams sharded across ranks, host gather, no collective"},
        "roofline": roof,
        "kernels": kb,
        "e2e": {"value": e2e_v, "unit": "frame-pairs/s",
                "h2d_bytes_per_step": fb + 4 * n, "d2h_bytes_per_step": 16 * n, "steps": e_steps,
                "api": "aof_px4_calc_flow_batch (pinned host frames -> H2D -> kernels -> D2H)" +
                       (", every step's records gathered on rank 0 through host shared memory (sharding.ShmGather), read one step behind" if world > 1 else ""),
                "roofline": {"bound": "pcie-h2d", "achieved": e2e_gbs, "peak": h2d_gbs, "unit": "GB/s per GPU",
                             "frac": e2e_gbs / h2d_gbs if h2d_gbs else None,
                             "peak_source": "measured in this run: pinned host -> device copy, all ranks at once, slowest rank"}},
        "gpu_launches": int(launches), "clocks": clocks, "dtype": "u8",
    }
    r["single_stream_latency_us"] = lat_us
    if multi is not None:
        r["e2e_host_threaded"] = multi
    if dense is not None:
        if "alu_packed_instr_per_s" in pk:
            a = dense["blocks_per_pair"] * 4944.0 * dense["pairs_per_s"]     # 289 x 16 SAD + 320 sub-pixel instr per block
            dense["roofline"] = {"bound": "int-alu", "achieved": a, "peak": pk["alu_packed_instr_per_s"], "unit": "packed-byte instr/s",
                                 "frac": a / pk["alu_packed_instr_per_s"], "algorithmic_instr_per_block": 4944,
                                 "kernel": "sad_dense_kernel<8>"}
        r["dense"] = dense
    if with_cpu:
        r["cpu_baseline"] = cpu_sad(frames, 10.0)
        r["vs_cpu"] = {"e2e_over_cpu": e2e_v / r["cpu_baseline"]["value"], "device_resident_over_cpu": r["value"] / r["cpu_baseline"]["value"],
                       "note": "GPU end-to-end (host frames in, results out) over the reference's CPU code on this box's %d host threads; "
                               "a reported baseline, not a quality measure" % r["cpu_baseline"]["cores"]}
    return r


def run_lk_gpu(args, rank, world, local, with_cpu):
    import torch
    pkg = entry.load_package()
    L = pkg.lib()
    n, rot = args.lk_streams, LK_ROT
    keep, h_frames = pinned_frames((rot, n, LK_H, LK_W))
    _, vel = lk_host_frames(n, rot, seed=2000 + rank, out=h_frames)
    fb = n * LK_W * LK_H
    d_frames = torch.empty((rot, fb), dtype=torch.uint8, device="cuda")
    d_frames.copy_(keep.reshape(rot, fb))
    d_t = torch.zeros((rot, n), dtype=torch.int32, device="cuda")
    d_out = torch.zeros((n, 4), dtype=torch.int32, device="cuda")
    d_raw = torch.zeros((n, 6), dtype=torch.int32, device="cuda")
    flow = pkg.LkFlow(FX, FY, -1, LK_W, LK_H, LK_FEATURES, n_streams=n, device=local)
    stream = torch.cuda.current_stream().cuda_stream
    fptr = [d_frames[k].data_ptr() for k in range(rot)]
    tptr = [d_t[k].data_ptr() for k in range(rot)]

    def step(k):
        flow.calc_flow_device(fptr[k % rot], tptr[k % rot], d_out.data_ptr(), d_raw.data_ptr(), stream=stream)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for k in range(args.warmup + 1):          # +1: the first frame only detects
        step(k)
    l0 = L.aof_launch_count()
    tm = time_steps(step, args.warmup + 1, args.steps, world)
    launches = (L.aof_launch_count() - l0) // tm["repeats"]
    ms, k = tm["ms"], tm["next"]
    clocks = sampler.stop(tm["t0"], tm["t1"]) if rank == 0 else None
    # parity spot check of the timed path: the flow of a translating texture is its translation
    raw = d_raw.cpu().numpy().view(pkg.capi.lk_raw_dtype).reshape(-1)
    p = (k - 1) % rot
    sign = 1.0 if 1 <= p <= rot // 2 else -1.0     # the last step went forward or back along the triangle wave
    ok = (np.abs(raw["px_x"] - sign * vel[:, 0]) < 1e-2) & (np.abs(raw["px_y"] - sign * vel[:, 1]) < 1e-2) & (raw["inliers"] >= 5)
    assert ok.mean() > 0.97, "timed LK path does not recover the translations (%.3f ok)" % ok.mean()
    bins_frac = float(raw["fast_bins"].mean()) / 16.0   # share of the 16 bins FAST ran on in the last timed step (steady state)
    lazy0 = flow.lazy_stats()
    kb = kernel_breakdown(pkg, step, k, min(args.steps, 50))
    lazy1 = flow.lazy_stats()
    k += min(args.steps, 50)

    # end to end through the host API: pinned host frames -> H2D -> kernels -> 16 B/stream D2H
    flow.close()
    flow_e = pkg.LkFlow(FX, FY, -1, LK_W, LK_H, LK_FEATURES, n_streams=n, device=local)
    t_us = np.zeros(n, np.uint32)
    for j in range(3):
        flow_e.calc_flow_batch(h_frames[j % rot], t_us + 1000 * j)
    e2e_ms, e_steps, res = time_e2e(lambda j: flow_e.calc_flow_batch(h_frames[(j + 3) % rot], t_us + 1000 * (j + 3)),
                                    0, min(args.steps, 30), world, rank)
    if rank == 0:
        assert res.size == world * n and (res["quality"] > 60).mean() > 0.97
    flow_e.close()
    h2d_gbs = h2d_probe(keep, world)
    multi = host_threaded_leg(
        lambda total, devs: pkg.LkFlowMulti(FX, FY, -1, LK_W, LK_H, LK_FEATURES, n_streams=total, devices=devs),
        h_frames, n, world, rank)

    # single-stream latency (BASELINE config 2 as written: one stream, host frame in, result out)
    one = pkg.LkFlow(FX, FY, -1, LK_W, LK_H, LK_FEATURES, n_streams=1, device=local)
    for j in range(4):
        one.calc_flow(h_frames[j % rot, 0], 1000 * j)
    t0 = time.perf_counter()
    for j in range(40):
        one.calc_flow(h_frames[(j + 4) % rot, 0], 1000 * (j + 4))
    lat_us = (time.perf_counter() - t0) / 40 * 1e6
    one.close()
    if rank != 0:
        return None
    pk = peaks()
    dom = max(kb, key=lambda kname: kb[kname]["ms_per_step"])
    step_ms = ms / args.steps
    # per-kernel rooflines.  Keys are prefixes of the kernel symbols the timing table reports.
    #   bytes: algorithmic HBM bytes per launch;  alu: packed integer instructions per launch
    px0 = LK_LEVEL_PX[0]
    lazy = [b1 - b0 for b0, b1 in zip(lazy0, lazy1)]
    k_steps = min(args.steps, 50)
    fast_alu = n * (px0 / 2.0) * bins_frac * LK_FAST_OPS_PER_PIXEL_PAIR
    spec = {
        "lk_pyr_stream": {"bound": "hbm", "bytes": n * LK_BYTES_PER_PAIR},          # level 0 read once + levels 1..3 written
        "lk_pyr_fused": {"bound": "hbm", "bytes": n * LK_BYTES_PER_PAIR},
        "lk_pyr_down": {"bound": "hbm", "bytes": n * (px0 + 2 * sum(LK_LEVEL_PX[1:3]) + sum(LK_LEVEL_PX[1:])) / 3.0},
        # cv::FAST of the reference reads every pixel of every bin that needs features once (trackFeatures.cpp:196): that is
        # the algorithmic traffic of the detector, however few of those pixels the kernel has to score
        "lk_fast_pick": {"bound": "hbm", "bytes": n * px0 * bins_frac},
        # the detector with every pixel scored: 72 packed u16x2 min / max per pixel pair on the ALU pipe
        "lk_fast_band": {"bound": "int-alu", "bytes": n * px0 * bins_frac, "alu": fast_alu},
        "lk_fast_kernel": {"bound": "int-alu", "bytes": n * px0 * bins_frac, "alu": fast_alu},
    }
    kroof = {}
    for kname, kv in kb.items():
        k_ms = kv["ms_per_step"] / max(1.0, kv["launches_per_step"])
        e = {"kernel_ms": k_ms, "bound": "latency"}
        sp = next((v for pre, v in spec.items() if kname.startswith(pre)), None)
        if sp and "flagged" not in kname:
            e["bound"] = sp["bound"]
            a = sp["bytes"] / (k_ms * 1e-3) / 1e9
            e["hbm"] = {"achieved": a, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": a / pk["hbm_gbs"],
                        "algorithmic_bytes_per_launch": sp["bytes"], "traffic": ncu_traffic(kname.split("<")[0], n, 512)}
            if "alu" in sp and "alu_packed_instr_per_s" in pk:
                a = sp["alu"] / (k_ms * 1e-3)
                e["int_alu"] = {"achieved": a, "peak": pk["alu_packed_instr_per_s"], "unit": "packed u16x2 min/max instr/s",
                                "bins_scored_frac": bins_frac, "frac": a / pk["alu_packed_instr_per_s"],
                                "algorithmic_instr_per_launch": sp["alu"],
                                "note": "algorithmic = the reference's cv::FAST: every pixel of every bin that needs features scored"}
        kroof[kname] = e
    step_bytes = n * LK_BYTES_PER_PAIR
    step_roof = {"algorithmic_bytes_per_step": step_bytes, "achieved": step_bytes / (step_ms * 1e-3) / 1e9,
                 "frac": step_bytes / (step_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                 "note": "the whole calcFlow step against HBM: new frame read once, pyramid levels 1..3 written (SURVEY.md 8(d))"}
    d = kroof[dom]
    if d["bound"] == "int-alu" and "int_alu" in d:
        roof = {"bound": "int-alu", "achieved": d["int_alu"]["achieved"], "peak": d["int_alu"]["peak"], "unit": d["int_alu"]["unit"],
                "frac": d["int_alu"]["frac"], "traffic": d["hbm"]["traffic"], "peak_source": pk.get("alu_which", ""),
                "algorithmic_instr_per_launch": d["int_alu"]["algorithmic_instr_per_launch"], "hbm": d["hbm"]}
    elif d["bound"] == "hbm":
        roof = {"bound": "hbm", "achieved": d["hbm"]["achieved"], "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": d["hbm"]["frac"],
                "traffic": d["hbm"]["traffic"], "peak_source": pk["which"], "algorithmic_bytes_per_launch": d["hbm"]["algorithmic_bytes_per_launch"]}
    else:
        # the kernel that takes the most time is latency-bound (per-feature iterations / serial selection): it has no
        # throughput roofline of its own; the step as a whole is stated against HBM instead
        roof = {"bound": "hbm", "achieved": step_roof["achieved"], "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": step_roof["frac"],
                "traffic": None, "peak_source": pk["which"], "algorithmic_bytes_per_launch": step_bytes,
                "note": "dominant kernel %s is latency-bound; this is the whole step against HBM" % dom}
    roof.update({"kernel": dom, "kernel_ms": d["kernel_ms"], "step": step_roof})
    e2e_v = world * n * e_steps / (e2e_ms * 1e-3)
    e2e_gbs = (fb + 4 * n) * e_steps / (e2e_ms * 1e-3) / 1e9
    r = {
        "value": world * n * args.steps / (ms * 1e-3), "ms_per_step": step_ms,
        "timed_region": {"repeats": tm["repeats"], "steps_per_repeat": args.steps, "total_s": sum(tm["all_ms"]) * 1e-3,
                         "ms_per_step_min": min(tm["all_ms"]) / args.steps, "ms_per_step_max": max(tm["all_ms"]) / args.steps},
        "config": {"workload": workload_name("lk", n),
                   "streams_per_gpu": n, "image": "640x480", "features": LK_FEATURES, "output_rate": -1,
                   "l2": "frames rotate over %d buffers = %d MiB > 126 MB L2" % (rot, rot * fb >> 20),
                   "parallelism": "streams sharded across ranks, host gather, no collective"},
        "roofline": roof,
        "kernels": kb, "kernel_rooflines": kroof,
        "fast": {"bins_needing_features_frac": bins_frac,
                 "detector": "masked band FAST (AOF_FAST_LAZY=2): only the pixels the first selection stage can accept from are scored; "
                             "the int-alu roofline counts the reference's work (every pixel of every bin that needs features)",
                 "per_step": {"bands_with_work": lazy[0] / k_steps, "live_units_16pairs_x_1row": lazy[1] / k_steps,
                              "live_unit_frac_of_needed_bins": (lazy[1] / k_steps) / max(1.0, n * 16 * bins_frac * 3 * 40 * 5),
                              "stream_frames_redone_on_the_complete_path": lazy[3] / k_steps}},
        "e2e": {"value": e2e_v, "unit": "frame-pairs/s",
                "h2d_bytes_per_step": fb + 4 * n, "d2h_bytes_per_step": (16 + 24) * n, "steps": e_steps,
                "api": "aof_lk_calc_flow_batch (pinned host frames -> H2D -> kernels -> D2H)" +
                       (", every step's records gathered on rank 0 through host shared memory (sharding.ShmGather), read one step behind" if world > 1 else ""),
                "roofline": {"bound": "pcie-h2d", "achieved": e2e_gbs, "peak": h2d_gbs, "unit": "GB/s per GPU",
                             "frac": e2e_gbs / h2d_gbs if h2d_gbs else None,
                             "peak_source": "measured in this run: pinned host -> device copy, all ranks at once, slowest rank"}},
        "single_stream_latency_us": lat_us,
        "e2e_host_threaded": multi,
        "overlap": "value: the device entry point pipelines two halves of the streams over two CUDA streams (the latency-bound "
                   "kernels of one half under the throughput kernels of the other); kernels / kernel_rooflines: the same steps with "
                   "per-kernel event timing on, which runs them unsplit, each kernel alone on the GPU (what a burst roofline "
                   "needs), so their sum exceeds ms_per_step; AOF_LK_SPLIT=1 times the unsplit step",
        "gpu_launches": int(launches), "clocks": clocks, "dtype": "u8 (i32 / f32 in LK)",
    }
    if with_cpu:
        r["cpu_baseline"] = cpu_lk(np.array(h_frames), 12.0)
        r["vs_cpu"] = {"e2e_over_cpu": e2e_v / r["cpu_baseline"]["value"], "device_resident_over_cpu": r["value"] / r["cpu_baseline"]["value"]}
    return r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="both", choices=["lk", "sad", "both"])
    ap.add_argument("--lk-streams", type=int, default=LK_STREAMS)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        if args.workload == "both":
            args.workload = "lk"
        run_reference(args)
        return
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        run_single_process(args)
        return
    rank, world, local = dist_setup()
    with_cpu = world == 1 and not args.no_cpu
    if with_cpu and rank == 0:
        entry.build_oracles()
    primary = "sad" if args.workload == "sad" else "lk"
    res = {}
    if args.workload in ("lk", "both"):
        res["lk"] = run_lk_gpu(args, rank, world, local, with_cpu)
    if args.workload in ("sad", "both"):
        sad_args = argparse.Namespace(**vars(args))
        sad_args.steps = args.steps * 10 if args.workload == "both" else args.steps   # SAD steps are ~100x shorter
        res["sad"] = run_sad_gpu(sad_args, rank, world, local, with_cpu)
        if res["sad"] is not None:
            res["sad"]["steps"] = sad_args.steps
        if world == 8:
            # BASELINE config 5 as written: 65 536 SAD streams over the 8 GPUs = 8192 per GPU (the weak-scaling series above
            # keeps config 3's 4096 per GPU at every N)
            c5 = run_sad_gpu(sad_args, rank, world, local, False, n_streams=SAD_STREAMS_CONFIG5, light=True)
            if c5 is not None:
                res["sad"]["config5"] = {k: c5[k] for k in ("value", "ms_per_step", "timed_region", "config", "roofline", "e2e", "gpu_launches")}
                res["sad"]["config5"]["total_streams"] = world * SAD_STREAMS_CONFIG5
    if rank != 0:
        return
    p = res[primary]
    line = {"metric": "frame-pairs/s", "value": p["value"], "unit": "frame-pairs/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": p["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": p["dtype"], "data": "synthetic", "config": p["config"], "roofline": p["roofline"],
            "e2e": p["e2e"], "gpu_launches": p["gpu_launches"], "clocks": p["clocks"], "kernels": p["kernels"]}
    if world > 1:
        import torch
        line["config"]["rank_to_device"] = ("local rank r -> device r // 2 + (r mod 2) * n / 2 of the n = {} visible devices (spread over the "
                                            "two host-side groups of the box, profiles/r2_h2d_scaling.json)").format(torch.cuda.device_count())
    for key in ("timed_region", "cpu_baseline", "vs_cpu", "single_stream_latency_us", "e2e_host_threaded", "kernel_rooflines", "fast", "dense", "overlap"):
        if key in p:
            line[key] = p[key]
    if args.workload == "both":
        s = res["sad"]
        line["sad"] = {k: s[k] for k in s if k not in ("clocks",)}
        line["sad"]["metric"], line["sad"]["unit"] = "frame-pairs/s", "frame-pairs/s"
    print(json.dumps(line))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
