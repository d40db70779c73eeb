"""aof_px4_calc_flow_pairs_static (csrc/aof_static.cu): the allocation-free entry point.  Its scratch is one fixed
__device__ array shared by every call, so the test makes two calls in a row (different data, different n, a rate-limited
and an unlimited configuration) and compares each, byte for byte (the bar of the other GPU-vs-GPU tests), with the
existing handle-based entry point aof_px4_calc_flow_device fed the same pairs.  Collected last on purpose (-x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
FX, FY = 216.6677, 216.2457


def _bind(L):
    L.aof_px4_static_max_pairs.restype = C.c_int
    L.aof_px4_calc_flow_pairs_static.restype = C.c_int
    L.aof_px4_calc_flow_pairs_static.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                                 C.c_void_p, C.c_void_p]


def _handle_path(pkg, prev, cur, t_us, rate, s):
    """the existing entry point: a fresh handle fed prev, then cur"""
    import torch
    torch.cuda.set_device(0)                                  # (earlier tests may have left another device current)
    n = prev.shape[0]
    f = pkg.Px4Flow(FX, FY, rate, 64, 64, s, 30, 3000, n_streams=n)
    d_prev, d_cur = torch.from_numpy(prev).to("cuda:0"), torch.from_numpy(cur).to("cuda:0")
    d_t0 = torch.zeros(n, dtype=torch.int32, device="cuda:0")
    d_t1 = torch.from_numpy(t_us.astype(np.int32)).to("cuda:0")
    d_out = torch.zeros((n, 4), dtype=torch.int32, device="cuda:0")
    f.calc_flow_device(d_prev.data_ptr(), d_t0.data_ptr(), d_out.data_ptr())
    f.calc_flow_device(d_cur.data_ptr(), d_t1.data_ptr(), d_out.data_ptr())
    torch.cuda.synchronize()
    out = d_out.cpu().numpy().view(pkg.capi.flow_out_dtype).reshape(-1).copy()
    f.close()
    return out


def _static_path(pkg, prev, cur, t_us, rate, s):
    import torch
    L = pkg.capi.lib()
    _bind(L)
    torch.cuda.set_device(0)
    n = prev.shape[0]
    p = pkg.capi.Px4Params(FX, FY, rate, 64, 64, s, 30, 3000)
    d_prev, d_cur = torch.from_numpy(prev).to("cuda:0"), torch.from_numpy(cur).to("cuda:0")
    d_t = torch.from_numpy(t_us.astype(np.int32)).to("cuda:0")
    d_out = torch.full((n, 4), -7, dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()                                  # everything is allocated BEFORE the call under test
    rc = L.aof_px4_calc_flow_pairs_static(C.byref(p), 0, d_prev.data_ptr(), d_cur.data_ptr(), d_t.data_ptr(), n,
                                          d_out.data_ptr(), None)
    assert rc == 0, L.aof_last_error()
    torch.cuda.synchronize()
    return d_out.cpu().numpy().view(pkg.capi.flow_out_dtype).reshape(-1).copy()


def test_static_entry_point_equals_handle_path_twice_in_a_row(pkg):
    rng = np.random.default_rng(77)
    # call 1: a large batch, rate-limited (the limiter sums of every pair are left in the shared array);
    # call 2: a smaller batch of OTHER pairs, other search size, unlimited -- must not see call 1's leftovers;
    # call 3: call 1's configuration again on new data
    cases = [(3000, 15, 4, 1), (517, -1, 6, 2), (3000, 15, 4, 3)]
    for n, rate, s, seed in cases:
        prev, cur, _ = pkg.synth.sad_pairs(n, max_shift=s, seed=seed, half_pixel_every=3)
        t_us = rng.integers(5000, 200000, n).astype(np.uint32)
        got = _static_path(pkg, prev, cur, t_us, rate, s)
        want = _handle_path(pkg, prev, cur, t_us, rate, s)
        assert got.tobytes() == want.tobytes(), (n, rate, s)
        assert (got["quality"] != -7).all()
    assert (want["quality"] > 0).mean() > 0.5                 # the comparison is not between two empty results


def test_static_entry_point_rejects_what_it_cannot_hold(pkg):
    import torch
    L = pkg.capi.lib()
    _bind(L)
    torch.cuda.set_device(0)
    cap = L.aof_px4_static_max_pairs()
    assert cap == 8192
    p = pkg.capi.Px4Params(FX, FY, -1, 64, 64, 4, 30, 3000)
    buf = torch.zeros(64 * 64 * 4 + 16, dtype=torch.uint8, device="cuda:0")
    t = torch.zeros(4, dtype=torch.int32, device="cuda:0")
    out = torch.zeros((4, 4), dtype=torch.int32, device="cuda:0")
    a = buf.data_ptr()
    assert L.aof_px4_calc_flow_pairs_static(C.byref(p), 0, a, a, t.data_ptr(), cap + 1, out.data_ptr(), None) != 0
    assert b"AOF_STATIC_MAX_PAIRS" in L.aof_last_error()
    assert L.aof_px4_calc_flow_pairs_static(C.byref(p), 0, a, a, t.data_ptr(), 0, out.data_ptr(), None) != 0
    assert L.aof_px4_calc_flow_pairs_static(C.byref(p), 0, a + 4, a, t.data_ptr(), 4, out.data_ptr(), None) != 0   # misaligned
    assert b"aligned" in L.aof_last_error()
    assert L.aof_px4_calc_flow_pairs_static(C.byref(p), 0, None, a, t.data_ptr(), 4, out.data_ptr(), None) != 0
    bad = pkg.capi.Px4Params(FX, FY, -1, 64, 64, 200, 30, 3000)
    assert L.aof_px4_calc_flow_pairs_static(C.byref(bad), 0, a, a, t.data_ptr(), 4, out.data_ptr(), None) != 0
