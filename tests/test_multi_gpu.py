"""aof_*_create_multi (aero-optical-flow_b200/csrc/aof_multi.cu): whole streams per device, one host worker thread per
device, one pinned result array as the host gather, no collective.

A multi handle over D devices must return, stream for stream, the bytes of ONE single-device handle over all streams
(streams are independent: src/flow_px4.cpp:69-98, src/flow_opencv.cpp:89-210 keep all state per instance), whatever the
split.  On a one-GPU box the D = 1 case runs (worker thread, pinned gather, ragged ranges are still exercised through
n_streams not divisible by anything); with 2+ devices every device count up to the box's is compared."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
FX, FY = 216.6677, 216.2457


def device_sets(pkg):
    n = pkg.capi.lib().aof_device_count()
    assert n >= 1
    sets = [[0]]
    if n >= 2:
        sets += [[0, 1], [1, 0]]
    if n >= 3:
        sets.append(list(range(n)))
    return sets


def test_px4_multi_equals_single_handle(pkg, port):
    s, T, n = 4, 4, 1003                                     # 1003 = 17 * 59: ragged over 2, 4 and 8 devices
    prev, cur, _ = pkg.synth.sad_pairs(n, max_shift=s, seed=21, half_pixel_every=3)
    frames = np.stack([prev, cur, prev, cur[::-1].copy()])
    rng = np.random.default_rng(4)
    t_us = np.stack([1000 + k * rng.integers(9000, 40000, n) for k in range(T)]).astype(np.uint32)
    for rate in (-1, 15):
        one = pkg.Px4Flow(FX, FY, rate, 64, 64, s, 30, 3000, n_streams=n)
        want = [one.calc_flow_batch(frames[k], t_us[k]).copy() for k in range(T)]
        one.close()
        for devs in device_sets(pkg):
            m = pkg.Px4FlowMulti(FX, FY, rate, 64, 64, s, 30, 3000, n_streams=n, devices=devs)
            rg = m.ranges()
            assert [r[0] for r in rg] == devs
            assert rg[0][1] == 0 and sum(r[2] for r in rg) == n
            assert all(rg[i][1] + rg[i][2] == rg[i + 1][1] for i in range(len(rg) - 1))      # contiguous whole-stream ranges
            assert max(r[2] for r in rg) - min(r[2] for r in rg) <= 1
            for k in range(T):
                got = m.calc_flow_batch(frames[k], t_us[k], copy=False)                     # the pinned gather array itself
                assert got.tobytes() == want[k].tobytes(), (rate, devs, k)
            m.close()
    # the seams against the oracle (first / last stream of every device's range at the largest split)
    devs = device_sets(pkg)[-1]
    m = pkg.Px4FlowMulti(FX, FY, -1, 64, 64, s, 30, 3000, n_streams=n, devices=devs)
    seams = sorted({i for _, f, c in m.ranges() for i in (f, f + c - 1)})
    got = [m.calc_flow_batch(frames[k], t_us[k]) for k in range(T)]
    m.close()
    for i in seams:
        o = port.px4(FX, FY, -1, 64, 64, s, 30, 3000)
        for k in range(T):
            q, dt, fx, fy, *_ = o.calc_flow(frames[k, i], int(t_us[k, i]))
            assert got[k]["quality"][i] == q, (i, k)
            if k and q >= 0:
                assert got[k]["dt_us"][i] == dt, (i, k)
                assert abs(got[k]["flow_x"][i] - fx) <= 2e-9 + 1.2e-7 * abs(fx)
                assert abs(got[k]["flow_y"][i] - fy) <= 2e-9 + 1.2e-7 * abs(fy)


def test_lk_multi_equals_single_handle(pkg):
    from oracle import lk as olk
    w, h, n, T, nf = 256, 256, 37, 4, 20
    rng = np.random.default_rng(12)
    vel = rng.integers(-3, 4, (n, 2))
    seqs = [pkg.synth.lk_sequence(T, w, h, int(vel[i, 0]), int(vel[i, 1]), seed=900 + i) for i in range(n)]
    frames = np.stack([np.stack([seqs[i][k] for i in range(n)]) for k in range(T)])
    t_us = np.stack([1000 + k * rng.integers(9000, 40000, n) for k in range(T)]).astype(np.uint32)
    for rate in (-1, 15):
        one = pkg.LkFlow(FX, FY, rate, w, h, nf, n_streams=n)
        want = [tuple(a.copy() for a in one.calc_flow_batch(frames[k], t_us[k], want_raw=True)) for k in range(T)]
        st_one = one.state()
        one.close()
        for devs in device_sets(pkg):
            m = pkg.LkFlowMulti(FX, FY, rate, w, h, nf, n_streams=n, devices=devs)
            for k in range(T):
                out, raw = m.calc_flow_batch(frames[k], t_us[k], want_raw=True)
                assert out.tobytes() == want[k][0].tobytes(), (rate, devs, k)
                assert raw.tobytes() == want[k][1].tobytes(), (rate, devs, k)
            st = m.state()
            for key in st_one:
                assert np.array_equal(st[key], st_one[key]), (devs, key)
            if rate == 15:                                                             # live setter on every device, state kept
                m.set_params(FX * 2, FY * 2, 15, 1.645)
                one2 = pkg.LkFlow(FX, FY, rate, w, h, nf, n_streams=n)
                for k in range(T):
                    one2.calc_flow_batch(frames[k], t_us[k])
                one2.set_params(FX * 2, FY * 2, 15, 1.645)
                t5 = t_us[T - 1] + 70000
                assert m.calc_flow_batch(frames[0], t5).tobytes() == one2.calc_flow_batch(frames[0], t5).tobytes()
                one2.close()
            m.close()
    devs = device_sets(pkg)[-1]
    m = pkg.LkFlowMulti(FX, FY, -1, w, h, nf, n_streams=n, devices=devs)
    seams = sorted({i for _, f, c in m.ranges() for i in (f, f + c - 1)})
    got = [m.calc_flow_batch(frames[k], t_us[k]) for k in range(T)]
    m.close()
    for i in seams:
        o = olk.Tracker(FX, FY, -1, w, h, nf)
        for k in range(T):
            q, dt, fx, fy = o.calc_flow(frames[k, i], int(t_us[k, i]))
            g = got[k]
            assert g["quality"][i] == q and g["flow_x"][i] == np.float32(fx) and g["flow_y"][i] == np.float32(fy), (i, k)


def test_multi_bad_arguments(pkg):
    L = pkg.capi.lib()
    p = pkg.capi.Px4Params(FX, FY, -1, 64, 64, 4, 30, 3000)
    h = C.c_void_p()
    devs = (C.c_int * 2)(0, 0)
    assert L.aof_px4_create_multi(C.byref(p), 8, devs, 2, C.byref(h)) != 0                # device listed twice
    assert b"twice" in L.aof_last_error()
    one = (C.c_int * 1)(0)
    assert L.aof_px4_create_multi(C.byref(p), 0, one, 1, C.byref(h)) != 0                 # fewer streams than devices
    bad = (C.c_int * 1)(L.aof_device_count() + 3)
    assert L.aof_px4_create_multi(C.byref(p), 8, bad, 1, C.byref(h)) != 0                 # no such device
    assert not h.value
