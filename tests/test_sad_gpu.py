"""GPU parity of the SAD path (K1 + epilogue) through the C ABI, bit-exact against the
oracle (oracle/sad_oracle.c, itself pinned to the reference) and the golden vectors."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "sad_golden.json")))
FX, FY = 216.6677, 216.2457


def angle(px, focal):
    return np.arctan2(np.asarray(px, np.float32).astype(np.float64), np.float64(np.float32(focal))).astype(np.float32)


def check_blocks(port, prev, cur, blocks, s, ft, vt):
    _, _, _, info = port.compute_flow(prev, cur, s, ft, vt, want_blocks=True)
    for k in range(25):
        e, g = info[k], blocks[k]
        assert (g["x"], g["y"]) == (e.x, e.y)
        assert g["texture"] == e.texture
        assert g["best_sad"] == e.best_sad, (k, g, e.best_sad)
        if e.searched:
            assert (g["dx"], g["dy"]) == (e.dx, e.dy)
        assert g["accepted"] == e.accepted and g["subdir"] == e.subdir
        if e.accepted:
            assert list(g["acc"]) == list(e.acc)


@pytest.mark.parametrize("case", GOLD["pairs"], ids=lambda c: "seed%d_s%d" % (c["seed"], c["search"]))
def test_pairs_vs_golden_and_oracle(pkg, port, case):
    s, vt = case["search"], case["val_thr"]
    prev, cur, _ = pkg.synth.sad_pairs(case["n"], max_shift=case["max_shift"], seed=case["seed"],
                                       half_pixel_every=case["half_pixel_every"])
    f = pkg.Px4Flow(FX, FY, -1, 64, 64, s, 30, vt, n_streams=1)
    raw, blocks = f.compute_flow_pairs(prev, cur, want_blocks=True)
    assert np.array_equal(raw["px_x"], np.array(case["fx"], np.float32))
    assert np.array_equal(raw["px_y"], np.array(case["fy"], np.float32))
    assert np.array_equal(raw["quality"], np.array(case["q"], np.int32))
    for i in range(0, case["n"], 7):   # per-block internals: SAD sums, displacements, acc[8], subdir
        check_blocks(port, prev[i], cur[i], blocks[i], s, 30, vt)
    f.close()


def test_kat_appendix_b2(pkg):
    tex = pkg.synth.lcg_texture()
    A = pkg.synth.crop(tex, 100, 100)
    B = pkg.synth.crop(tex, 98, 101)
    f = pkg.Px4Flow(FX, FY, -1, 64, 64, 4, 30, 3000)
    raw, blocks = f.compute_flow_pairs(A[None], B[None], want_blocks=True)
    b0 = blocks[0][0]
    assert (b0["x"], b0["y"], b0["texture"], b0["best_sad"], b0["dx"], b0["dy"]) == (5, 5, 340, 0, 2, -1)
    assert list(b0["acc"]) == [549, 681, 411, 673, 527, 617, 427, 656] and b0["subdir"] == 8
    assert (raw["px_x"][0], raw["px_y"][0], raw["quality"][0], raw["count"][0]) == (2.0, -1.0, 255, 25)
    for e in GOLD["kat"]["lcg_shifts"]:
        if e["s"] != 4:
            continue
        Bs = pkg.synth.crop(tex, 100 - e["dx"], 100 - e["dy"])
        r = f.compute_flow_pairs(A[None], Bs[None])
        assert (r["px_x"][0], r["px_y"][0], r["quality"][0]) == (np.float32(e["fx"]), np.float32(e["fy"]), e["q"])
    Bh = ((B.astype(np.int32) + pkg.synth.crop(tex, 97, 101) + 1) // 2).astype(np.uint8)
    r = f.compute_flow_pairs(A[None], Bh[None])
    assert (r["px_x"][0], r["px_y"][0], r["quality"][0]) == (2.5, -1.0, 255)
    f.close()


def test_edge_inputs(pkg, port):
    """low texture, uncorrelated noise, flat, saturated, gate/accept thresholds at the boundary."""
    rng = np.random.default_rng(3)
    frames = [np.full((64, 64), 0, np.uint8), np.full((64, 64), 255, np.uint8),
              (128 + rng.integers(0, 3, (64, 64))).astype(np.uint8),
              rng.integers(0, 256, (64, 64), dtype=np.uint8), rng.integers(0, 256, (64, 64), dtype=np.uint8),
              np.tile(np.arange(64, dtype=np.uint8) * 4, (64, 1)), np.tile(np.arange(64, dtype=np.uint8)[:, None] * 4, (1, 64))]
    prev = np.stack([a for a in frames for _ in frames])
    cur = np.stack([b for _ in frames for b in frames])
    for s, ft, vt in [(4, 30, 3000), (6, 0, 20000), (8, 30, 3000), (4, 340, 1), (4, 341, 0xFFFFFFF)]:
        f = pkg.Px4Flow(FX, FY, -1, 64, 64, s, ft, vt)
        raw, blocks = f.compute_flow_pairs(prev, cur, want_blocks=True)
        ex, ey, eq = port.compute_flow_batch(prev, cur, s, ft, vt)
        assert np.array_equal(raw["px_x"], ex) and np.array_equal(raw["px_y"], ey) and np.array_equal(raw["quality"], eq)
        for i in range(len(prev)):
            check_blocks(port, prev[i], cur[i], blocks[i], s, ft, vt)
        f.close()


@pytest.mark.parametrize("w,h", [(32, 32), (48, 64), (96, 96), (128, 128), (252, 256)])
def test_other_geometries(pkg, port, w, h):
    rng = np.random.default_rng(w)
    n = 12
    base = pkg.synth.box_blur3(rng.integers(0, 256, (n, h + 16, w + 16), dtype=np.uint8))
    prev = np.ascontiguousarray(base[:, 8:8 + h, 8:8 + w])
    for s in (4, 6, 8):
        if w < 2 * (s + 1) + 9:
            continue
        sh = rng.integers(-s, s + 1, (n, 2))
        cur = np.stack([base[i, 8 - sh[i, 1]:8 - sh[i, 1] + h, 8 - sh[i, 0]:8 - sh[i, 0] + w] for i in range(n)])
        f = pkg.Px4Flow(FX, FY, -1, w, h, s, 30, 3000)
        raw = f.compute_flow_pairs(prev, cur)
        for i in range(n):
            ex, ey, eq = port.compute_flow(prev[i], cur[i], s, 30, 3000)
            assert (raw["px_x"][i], raw["px_y"][i], raw["quality"][i]) == (np.float32(ex), np.float32(ey), eq)
        f.close()


@pytest.mark.parametrize("seq", GOLD["seq"], ids=lambda s: "rate%d" % s["rate"])
def test_calcflow_sequence_single_stream(pkg, seq):
    """aof_px4_calc_flow == OpticalFlowPX4::calcFlow on the reference's recorded sequence."""
    tex = pkg.synth.lcg_texture()
    f = pkg.Px4Flow(FX, FY, seq["rate"], 64, 64, 4, 30, 3000)
    for t, row in enumerate(seq["rows"]):
        q, dt, fx, fy = f.calc_flow(pkg.synth.crop(tex, 100 - 2 * t, 100 + t), 1000 + seq["dt_step"] * t)
        assert q == row["q"]
        if row["fx"] is None:
            assert np.isnan(fx) and np.isnan(fy) and dt == -123456
        else:
            # double atan2 on device vs glibc: equal up to one float ulp (<< the 1e-3 px budget)
            assert abs(fx - row["fx"]) <= 2e-9 + 1.2e-7 * abs(row["fx"])
            assert abs(fy - row["fy"]) <= 2e-9 + 1.2e-7 * abs(row["fy"])
        assert dt == row["dt"]
    f.close()


def test_batch_streams_equal_independent_oracle_runs(pkg, port):
    """N streams in one handle == N single-stream oracle objects, for any batch order; rate limiter on."""
    n, T, s = 96, 9, 4
    rng = np.random.default_rng(8)
    tex = pkg.synth.box_blur3(rng.integers(0, 256, (n, 64 + 80, 64 + 80), dtype=np.uint8))
    vel = rng.integers(-4, 5, (n, 2))
    frames = np.stack([np.stack([tex[i, 40 - k * vel[i, 1]:104 - k * vel[i, 1], 40 - k * vel[i, 0]:104 - k * vel[i, 0]]
                                 for i in range(n)]) for k in range(T)])          # [T, n, 64, 64]
    tstep = rng.integers(9000, 30000, n).astype(np.uint32)
    perm = rng.permutation(n)
    for rate in (-1, 15):
        f = pkg.Px4Flow(FX, FY, rate, 64, 64, s, 30, 3000, n_streams=n)
        g = pkg.Px4Flow(FX, FY, rate, 64, 64, s, 30, 3000, n_streams=n)      # same streams, permuted
        orc = [port.px4(FX, FY, rate, 64, 64, s, 30, 3000) for _ in range(n)]
        for k in range(T):
            t_us = (1000 + tstep * k).astype(np.uint32)
            out = f.calc_flow_batch(frames[k], t_us)
            outp = g.calc_flow_batch(frames[k][perm], t_us[perm])
            assert np.array_equal(out[perm], outp)
            for i in range(n):
                q, dt, fx, fy, *_ = orc[i].calc_flow(frames[k, i], int(t_us[i]))
                assert out["quality"][i] == q
                if k == 0:
                    continue
                assert abs(out["flow_x"][i] - fx) <= 2e-9 + 1.2e-7 * abs(fx)
                assert abs(out["flow_y"][i] - fy) <= 2e-9 + 1.2e-7 * abs(fy)
                if q >= 0 or rate <= 0:
                    assert out["dt_us"][i] == dt
        f.close(); g.close()


def test_full_size_batch_property(pkg):
    """BASELINE config 3 size (4096 streams): integer translations within +-s are recovered exactly."""
    n, s = 4096, 4
    prev, cur, sh = pkg.synth.sad_pairs(n, max_shift=s, seed=21)
    f = pkg.Px4Flow(FX, FY, -1, 64, 64, s, 30, 3000, n_streams=n)
    t = np.zeros(n, np.uint32)
    f.calc_flow_batch(prev, t)
    out = f.calc_flow_batch(cur, t + 5000)
    assert (out["quality"] == 255).all() and (out["dt_us"] == 5000).all()
    assert np.array_equal(out["flow_x"], angle(sh[:, 0], FX)) or np.allclose(out["flow_x"], angle(sh[:, 0], FX), rtol=1.2e-7, atol=2e-9)
    assert np.allclose(out["flow_y"], angle(sh[:, 1], FY), rtol=1.2e-7, atol=2e-9)
    # swapping the frames negates the flow (antisymmetry of pure translation)
    f.reset()
    f.calc_flow_batch(cur, t)
    back = f.calc_flow_batch(prev, t + 5000)
    assert np.allclose(back["flow_x"], -out["flow_x"], atol=1e-7) and np.allclose(back["flow_y"], -out["flow_y"], atol=1e-7)
    f.close()


@pytest.mark.parametrize("s", [1, 2, 3, 5, 7, 10, 26])
def test_any_search_size_vs_reference_objects(pkg, port, ref, s):
    """flow_px4.hpp:62-65 takes any search size: sizes without a specialised kernel run sad_flow_any_kernel and
    must equal the reference objects (and the port, block by block)."""
    prev, cur, _ = pkg.synth.sad_pairs(40, max_shift=min(s, 8), seed=300 + s, half_pixel_every=3)
    f = pkg.Px4Flow(FX, FY, -1, 64, 64, s, 30, 3000, n_streams=1)
    raw, blocks = f.compute_flow_pairs(prev, cur, want_blocks=True)
    rx, ry, rq = ref.compute_flow_batch(prev, cur, s, 30, 3000)
    assert np.array_equal(raw["px_x"], rx) and np.array_equal(raw["px_y"], ry) and np.array_equal(raw["quality"], rq)
    if s <= 10:
        assert (rq == 255).sum() >= 30          # the run is not vacuous
    else:
        assert (rq == 0).all()                  # 64 - 2 * 27 - 8 = 2 px of grid: one position, never > 10 votes
    if s <= 10:                                 # a full 5 x 5 grid: every block record is comparable
        for i in (0, 17, 39):
            check_blocks(port, prev[i], cur[i], blocks[i], s, 30, 3000)
    f.close()


def test_misaligned_device_pointer_is_rejected(pkg):
    """the device entry points stage frames with bulk (TMA-engine) copies: a source that is not 16-byte aligned is
    refused with AOF_ERR_INVALID instead of faulting the context (include/aof.h)"""
    import torch
    n = 8
    buf = torch.zeros(n * 4096 * 2 + 64, dtype=torch.uint8, device="cuda")
    t = torch.zeros(n, dtype=torch.int32, device="cuda")
    out = torch.zeros(n * 4, dtype=torch.int32, device="cuda")
    raw = torch.zeros(n * 4, dtype=torch.int32, device="cuda")
    f = pkg.Px4Flow(FX, FY, -1, 64, 64, 4, 30, 3000, n_streams=n)
    L = pkg.lib()
    assert L.aof_px4_calc_flow_device(f.handle, buf.data_ptr() + 1, t.data_ptr(), out.data_ptr(), 0, 0, 0) == -100
    assert b"16-byte" in L.aof_last_error()
    assert L.aof_px4_compute_flow_pairs_device(f.handle, buf.data_ptr() + 4, buf.data_ptr() + n * 4096, n, raw.data_ptr(), 0, 0) == -100
    assert L.aof_px4_compute_flow_pairs_device(f.handle, buf.data_ptr(), buf.data_ptr() + n * 4096 + 8, n, raw.data_ptr(), 0, 0) == -100
    # the context is still healthy and an aligned call works
    assert L.aof_px4_compute_flow_pairs_device(f.handle, buf.data_ptr(), buf.data_ptr() + n * 4096, n, raw.data_ptr(), 0, 0) == 0
    torch.cuda.synchronize()
    f.close()
