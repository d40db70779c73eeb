"""The host-batch entry points cut a large batch into chunks of whole streams and overlap the upload of chunk c + 1 with the
kernels of chunk c (aof_px4.cu / aof_lk.cu).  Stream counts that do not divide into the chunks, pageable and page-locked
callers' buffers, and the device-resident entry point (one launch over all streams, no chunks) must all give the same
bytes; a sample of streams around the chunk seams is checked against the oracle."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
FX, FY = 216.6677, 216.2457


def device_run_px4(pkg, frames, t_us, rate, s):
    import torch
    T, n, h, w = frames.shape
    f = pkg.Px4Flow(FX, FY, rate, w, h, s, 30, 3000, n_streams=n)
    d_out = torch.zeros((n, 4), dtype=torch.int32, device="cuda")
    outs = []
    for k in range(T):
        d_img = torch.from_numpy(frames[k]).cuda()
        d_t = torch.from_numpy(t_us[k].astype(np.int32)).cuda()
        torch.cuda.synchronize()
        f.calc_flow_device(d_img.data_ptr(), d_t.data_ptr(), d_out.data_ptr())
        torch.cuda.synchronize()
        outs.append(d_out.cpu().numpy().view(pkg.capi.flow_out_dtype).reshape(-1).copy())
        keep = d_img                                                             # the handle reads it as `prev` next call
    f.close()
    return outs


@pytest.mark.parametrize("n", [1111, 2049])
def test_px4_chunked_batch_equals_device_path_and_oracle(pkg, port, n):
    s, T = 4, 4
    prev, cur, _ = pkg.synth.sad_pairs(n, max_shift=s, seed=5, half_pixel_every=3)
    rng = np.random.default_rng(3)
    frames = np.stack([prev, cur, prev, cur[::-1].copy()])                       # [T, n, 64, 64]
    assert frames[0].nbytes >> 21 >= 2                                           # really more than one chunk
    t_us = np.stack([1000 + k * rng.integers(9000, 40000, n) for k in range(T)]).astype(np.uint32)
    for rate in (-1, 15):
        f = pkg.Px4Flow(FX, FY, rate, 64, 64, s, 30, 3000, n_streams=n)
        got = [f.calc_flow_batch(frames[k], t_us[k]).copy() for k in range(T)]
        f.close()
        # page-locked caller buffer: the copy stream reads it in place
        L = pkg.capi.lib()
        pinned = frames.copy()
        assert L.aof_host_register(C.c_void_p(pinned.ctypes.data), C.c_size_t(pinned.nbytes)) == 0
        try:
            g = pkg.Px4Flow(FX, FY, rate, 64, 64, s, 30, 3000, n_streams=n)
            got_pinned = [g.calc_flow_batch(pinned[k], t_us[k]).copy() for k in range(T)]
            g.close()
        finally:
            assert L.aof_host_unregister(C.c_void_p(pinned.ctypes.data)) == 0
        want = device_run_px4(pkg, frames, t_us, rate, s)
        for k in range(T):
            assert got[k].tobytes() == got_pinned[k].tobytes(), (rate, k)
            for fld in ("quality", "flow_x", "flow_y"):
                assert np.array_equal(got[k][fld], want[k][fld]), (rate, k, fld)
            ok = (got[k]["quality"] >= 0) | (rate <= 0)
            assert np.array_equal(got[k]["dt_us"][ok], want[k]["dt_us"][ok]), (rate, k)
        # oracle on the streams either side of every chunk seam
        chunks = min(frames[0].nbytes >> 21, 8)
        per = (n + chunks - 1) // chunks
        sample = sorted({0, n - 1} | {i for c in range(1, chunks) for i in (c * per - 1, c * per) if i < n})
        for i in sample:
            o = port.px4(FX, FY, rate, 64, 64, s, 30, 3000)
            for k in range(T):
                q, dt, fx, fy, *_ = o.calc_flow(frames[k, i], int(t_us[k, i]))
                assert got[k]["quality"][i] == q, (rate, k, i)
                if k and q >= 0:
                    assert abs(got[k]["flow_x"][i] - fx) <= 2e-9 + 1.2e-7 * abs(fx)
                    assert abs(got[k]["flow_y"][i] - fy) <= 2e-9 + 1.2e-7 * abs(fy)
                    assert got[k]["dt_us"][i] == dt


@pytest.mark.parametrize("split", ["1", "2"])
def test_lk_chunked_batch_equals_device_path_and_oracle(pkg, split, monkeypatch):
    """split = 2: the device entry point runs the two halves of the streams on two CUDA streams (fork / join by events)."""
    import torch
    monkeypatch.setenv("AOF_LK_SPLIT", split)
    from oracle import lk as olk
    w, h, n, T, nf = 256, 256, 50, 4, 20
    rng = np.random.default_rng(11)
    vel = rng.integers(-3, 4, (n, 2))
    seqs = [pkg.synth.lk_sequence(T, w, h, int(vel[i, 0]), int(vel[i, 1]), seed=300 + i) for i in range(n)]
    frames = np.stack([np.stack([seqs[i][k] for i in range(n)]) for k in range(T)])      # [T, n, h, w]
    chunks = min(n // 16, frames[0].nbytes >> 20, 8)
    assert chunks == 3
    t_us = np.stack([1000 + k * rng.integers(9000, 40000, n) for k in range(T)]).astype(np.uint32)
    for rate in (-1, 15):
        f = pkg.LkFlow(FX, FY, rate, w, h, nf, n_streams=n)
        got = [tuple(a.copy() for a in f.calc_flow_batch(frames[k], t_us[k], want_raw=True)) for k in range(T)]
        st_host = f.state()
        f.close()
        g = pkg.LkFlow(FX, FY, rate, w, h, nf, n_streams=n)
        d_out = torch.zeros((n, 4), dtype=torch.int32, device="cuda")
        d_raw = torch.zeros(n * pkg.capi.lk_raw_dtype.itemsize, dtype=torch.uint8, device="cuda")
        for k in range(T):
            d_img = torch.from_numpy(frames[k]).cuda()
            d_t = torch.from_numpy(t_us[k].astype(np.int32)).cuda()
            torch.cuda.synchronize()
            g.calc_flow_device(d_img.data_ptr(), d_t.data_ptr(), d_out.data_ptr(), d_raw.data_ptr())
            torch.cuda.synchronize()
            out = d_out.cpu().numpy().view(pkg.capi.flow_out_dtype).reshape(-1)
            raw = d_raw.cpu().numpy().view(pkg.capi.lk_raw_dtype).reshape(-1)
            assert got[k][0].tobytes() == out.tobytes(), (rate, k)
            assert got[k][1].tobytes() == raw.tobytes(), (rate, k)
            keep = d_img                                                         # level 0 of the previous pyramid next call
        st_dev = g.state()
        g.close()
        for key in st_host:
            assert np.array_equal(st_host[key], st_dev[key]), key
        per = (n + chunks - 1) // chunks
        for i in sorted({0, n - 1} | {j for c in range(1, chunks) for j in (c * per - 1, c * per)}):
            o = olk.Tracker(FX, FY, rate, w, h, nf)
            for k in range(T):
                q, dt, fx, fy = o.calc_flow(frames[k, i], int(t_us[k, i]))
                out = got[k][0]
                assert out["quality"][i] == q and out["flow_x"][i] == np.float32(fx) and out["flow_y"][i] == np.float32(fy), (rate, k, i)
                if q >= 0 or rate <= 0:
                    assert out["dt_us"][i] == dt
