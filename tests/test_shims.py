"""The header-compatible C++ classes (aero-optical-flow_b200/cpp) compile against include/aof.h with a
C++11 compiler (the reference's standard) and, on the GPU, reproduce the oracle call by call."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "aero-optical-flow_b200")
EXE = os.path.join(ROOT, "tests", "_shim_main")
FX, FY = 216.6677, 216.2457


def build_driver(pkg):
    cmd = ["g++", "-std=c++11", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(PKG, "cpp"),
           os.path.join(ROOT, "tests", "shim_main.cpp"), "-o", EXE, "-L" + PKG, "-laof_b200", "-Wl,-rpath," + PKG]
    subprocess.run(cmd, check=True)
    return EXE


def test_shims_compile_and_link(pkg):
    assert os.path.exists(build_driver(pkg))
    # without a GPU every call fails loudly and calcFlow answers -1 (never a CPU fallback)
    import torch
    if not torch.cuda.is_available():
        raw = os.path.join(ROOT, "tests", "_shim_frames.raw")
        np.zeros((2, 64, 64), np.uint8).tofile(raw)
        r = subprocess.run([EXE, "px4", raw, "64", "64", "2", "-1", "10000"], capture_output=True, text=True)
        assert r.returncode == 0 and all(l.split()[0] == "-1" for l in r.stdout.strip().splitlines())
        assert "aof_px4_create failed" in r.stderr
        os.remove(raw)


def run_driver(mode, frames, rate, step):
    raw = os.path.join(ROOT, "tests", "_shim_frames.raw")
    frames.tofile(raw)
    T, h, w = frames.shape
    r = subprocess.run([EXE, mode, raw, str(w), str(h), str(T), str(rate), str(step)], capture_output=True, text=True, check=True)
    os.remove(raw)
    rows = []
    for line in r.stdout.strip().splitlines():
        q, dt, fx, fy = line.split()
        rows.append((int(q), int(dt), float.fromhex(fx), float.fromhex(fy)))
    return rows


@pytest.mark.gpu
@pytest.mark.parametrize("rate", [-1, 15])
def test_px4_class_matches_oracle(pkg, port, rate):
    build_driver(pkg)
    tex = pkg.synth.lcg_texture()
    frames = np.stack([pkg.synth.crop(tex, 100 - 2 * t, 100 + t) for t in range(10)])
    rows = run_driver("px4", frames, rate, 13333)
    o = port.px4(FX, FY, rate, 64, 64, 4, 30, 3000)
    for k, (q, dt, fx, fy) in enumerate(rows):
        oq, odt, ofx, ofy, *_ = o.calc_flow(frames[k], 1000 + 13333 * k)
        assert q == oq
        if k == 0:
            assert (dt, fx, fy) == (-123456, -77.0, -77.0)      # first call leaves the outputs untouched
            continue
        assert abs(fx - ofx) <= 2e-9 + 1.2e-7 * abs(ofx) and abs(fy - ofy) <= 2e-9 + 1.2e-7 * abs(ofy)
        assert dt == (odt if (oq >= 0 or rate <= 0) else -123456)


@pytest.mark.gpu
@pytest.mark.parametrize("mode,rate", [("lk", -1), ("lk", 15), ("lkcam", -1)])
def test_opencv_class_matches_oracle(pkg, mode, rate):
    from oracle import lk as olk
    build_driver(pkg)
    frames = np.stack(pkg.synth.lk_sequence(8, 128, 128, 2, -1, seed=31))
    rows = run_driver(mode, frames, rate, 13333)
    o = olk.Tracker(FX, FY, rate, 128, 128, 20)
    if mode == "lkcam":
        o.set_camera_matrix(FX, FY, 64.0, 64.0)
        o.set_camera_distortion(-0.2, 0.05, 0.01, 0.001, -0.002)
    for k, (q, dt, fx, fy) in enumerate(rows):
        oq, odt, ofx, ofy = o.calc_flow(frames[k], 1000 + 13333 * k)
        assert (q, fx, fy) == (oq, ofx, ofy), k
        assert dt == (odt if (oq >= 0 or rate <= 0) else -123456)


RETUNE = {4: (300.0, 310.0, 15, 1.645), 8: (300.0, 310.0, -1, 1.0), 11: (216.0, 216.0, 20, 2.5)}


@pytest.mark.gpu
def test_px4_setters_keep_state(pkg, port):
    """setFocalLength* / setOutputRate between frames: the reference's setters only assign
    (include/optical_flow.hpp:74-78); previous frame, limiter sums and time_last_pub survive."""
    build_driver(pkg)
    tex = pkg.synth.lcg_texture()
    frames = np.stack([pkg.synth.crop(tex, 100 - 2 * t, 100 + t) for t in range(14)])
    rows = run_driver("px4retune", frames, 15, 13333)
    o = port.px4(FX, FY, 15, 64, 64, 4, 30, 3000)
    rate = 15
    for k, (q, dt, fx, fy) in enumerate(rows):
        if k in RETUNE:
            o.set_params(*RETUNE[k][:3])
            rate = RETUNE[k][2]
        oq, odt, ofx, ofy, *_ = o.calc_flow(frames[k], 1000 + 13333 * k)
        assert q == oq, k
        if k == 0:
            continue
        assert abs(fx - ofx) <= 2e-9 + 1.2e-7 * abs(ofx) and abs(fy - ofy) <= 2e-9 + 1.2e-7 * abs(ofy)
        assert dt == (odt if (oq >= 0 or rate <= 0) else -123456), k


@pytest.mark.gpu
def test_opencv_setters_keep_state(pkg):
    from oracle import lk as olk
    build_driver(pkg)
    frames = np.stack(pkg.synth.lk_sequence(14, 128, 128, 1, 1, seed=34))
    rows = run_driver("lkretune", frames, 15, 13333)
    o = olk.Tracker(FX, FY, 15, 128, 128, 20)
    rate = 15
    for k, (q, dt, fx, fy) in enumerate(rows):
        if k in RETUNE:
            o.set_params(*RETUNE[k])
            rate = RETUNE[k][2]
        oq, odt, ofx, ofy = o.calc_flow(frames[k], 1000 + 13333 * k)
        assert (q, fx, fy) == (oq, ofx, ofy), k
        assert dt == (odt if (oq >= 0 or rate <= 0) else -123456)
