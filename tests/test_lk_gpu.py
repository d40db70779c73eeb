"""GPU parity of the LK path (K3 pyramid, K6 LK, K4 FAST, K5 selection, K7 epilogue) through the
C ABI against oracle/lk_oracle.c (pinned to cv2 4.13 by tests/test_oracle_lk.py) and the golden
vectors.  Integer work (pyramid, FAST scores, status, update vectors, quality) is bit-exact; LK
positions and flow are bit-exact against the oracle and within 2e-4 px / 1e-3 px of cv2."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import lk as olk

pytestmark = pytest.mark.gpu
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "lk_golden.json")))
FX, FY = 216.6677, 216.2457
ANG_TOL = 1e-3 / 216.0


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


@pytest.mark.parametrize("w,h", [(640, 480), (128, 128), (64, 64), (75, 61), (322, 242)])
def test_pyramid_bit_exact(pkg, w, h):
    imgs = np.stack([pkg.synth.lk_sequence(1, w, h, 1, 1, seed=s)[0] for s in range(3)])
    f = pkg.LkFlow(FX, FY, -1, w, h)
    levels = f.pyramid(imgs)
    assert len(levels) == olk.pyr_top_level(w, h) + 1
    for i in range(len(imgs)):
        ref = olk.build_pyramid(imgs[i])
        for l in range(len(levels)):
            assert np.array_equal(levels[l][i], ref[l]), (i, l)
    f.close()


@pytest.mark.parametrize("w,h,bands", [(640, 480, 1), (640, 480, 2), (640, 480, 7), (640, 480, 60), (1280, 720, 3), (96, 80, 1),
                                       (256, 88, 2), (288, 80, 1), (544, 128, 4)])
def test_pyramid_stream_kernel_bands(pkg, w, h, bands, monkeypatch):
    """the streaming 3-level kernel (W % 32 == 0, H % 8 == 0) with forced band counts: every strip / band seam and
    every per-level reflect-101 border against the oracle"""
    monkeypatch.setenv("AOF_PYR_BANDS", str(bands))
    rng = np.random.default_rng(w * 7 + h + bands)
    imgs = rng.integers(0, 256, (2, h, w), dtype=np.uint8)
    f = pkg.LkFlow(FX, FY, -1, w, h)
    levels = f.pyramid(imgs)
    assert len(levels) == 4
    for i in range(len(imgs)):
        ref = olk.build_pyramid(imgs[i])
        for l in range(4):
            assert np.array_equal(levels[l][i], ref[l]), (i, l, np.argwhere(levels[l][i] != ref[l])[:5])
    f.close()


def test_pyramid_golden(pkg):
    for case in GOLD["pyr"]:
        img = pkg.synth.lk_sequence(1, case["w"], case["h"], 1, 1, seed=case["seed"])[0]
        f = pkg.LkFlow(FX, FY, -1, case["w"], case["h"])
        levels = f.pyramid(img[None])
        for l in range(1, len(levels)):
            assert sha(levels[l][0]) == case["levels"][l - 1]
        f.close()


@pytest.mark.parametrize("w,h", [(640, 480), (128, 128), (64, 64), (100, 76), (1280, 720), (352, 300), (2048, 160)])
def test_fast_scores_bit_exact(pkg, w, h):
    imgs = np.stack([pkg.synth.lk_sequence(1, w, h, 1, 1, seed=10 + s)[0] for s in range(2)])
    rng = np.random.default_rng(w)
    imgs = np.concatenate([imgs, rng.integers(0, 256, (1, h, w), dtype=np.uint8)])    # raw noise: dense corners, many ties
    f = pkg.LkFlow(FX, FY, -1, w, h)
    sc = f.fast_scores(imgs)
    bw, bh = w // 4, h // 4
    for i in range(len(imgs)):
        want = np.zeros((h, w), np.uint8)
        for bx in range(4):
            for by in range(4):
                want[by * bh:(by + 1) * bh, bx * bw:(bx + 1) * bw] = olk.fast(imgs[i][by * bh:(by + 1) * bh, bx * bw:(bx + 1) * bw])
        assert np.array_equal(sc[i], want), i
    f.close()


@pytest.mark.parametrize("case", GOLD["lk"], ids=lambda c: "%dx%d" % (c["w"], c["h"]))
def test_track_points_golden_and_oracle(pkg, case):
    """LK with the corner set injected (north_star: 'after matching the selected corner set')."""
    fr = pkg.synth.lk_sequence(2, case["w"], case["h"], case["dx"], case["dy"], seed=case["seed"])
    pts = np.array(case["pts"], np.float32)
    f = pkg.LkFlow(FX, FY, -1, case["w"], case["h"])
    out, st = f.track_points(fr[0][None], fr[1][None], pts[None])
    o_out, o_st = olk.lk(fr[0], fr[1], pts)
    assert np.array_equal(st[0], o_st) and np.array_equal(st[0], np.array(case["status"], np.uint8))
    ok = o_st == 1
    assert np.array_equal(out[0][ok], o_out[ok])                                  # bit-exact vs the oracle
    assert np.abs(out[0][ok] - np.array(case["next"])[ok]).max() <= 2e-4            # vs cv2
    f.close()


def test_track_points_batch_borders(pkg):
    """many pairs x many points, incl. points on/over the border and untrackable (flat) frames."""
    w, h, n = 160, 120, 6
    rng = np.random.default_rng(3)
    prev, cur = [], []
    for i in range(n):
        fr = pkg.synth.lk_sequence(2, w, h, int(rng.integers(-6, 7)), int(rng.integers(-6, 7)), seed=40 + i)
        prev.append(fr[0]); cur.append(fr[1])
    prev[-1] = np.full((h, w), 90, np.uint8)
    prev, cur = np.stack(prev), np.stack(cur)
    pts = rng.uniform([-12, -12], [w + 12, h + 12], (n, 64, 2)).astype(np.float32)
    pts[:, :4] = [[0, 0], [w - 1, h - 1], [w, h], [w / 2, -4.5]]
    f = pkg.LkFlow(FX, FY, -1, w, h)
    out, st = f.track_points(prev, cur, pts)
    for i in range(n):
        o_out, o_st = olk.lk(prev[i], cur[i], pts[i])
        assert np.array_equal(st[i], o_st), i
        assert np.array_equal(out[i][o_st == 1], o_out[o_st == 1]), i
    f.close()


def test_track_points_final_position_retest(pkg):
    """the level-0 re-test of the final window origin (cv2 known answer, tests/test_oracle_lk.py) + the band around it"""
    fr = pkg.synth.lk_sequence(2, 160, 120, 3, 0, seed=75)
    rng = np.random.default_rng(5)
    pts = np.stack([161 + rng.uniform(-0.3, 0.3, 2000), rng.uniform(10, 110, 2000)], 1).astype(np.float32)
    pts[0] = [160.72454833984375, 53.756690979003906]
    f = pkg.LkFlow(FX, FY, -1, 160, 120)
    out, st = f.track_points(fr[0][None], fr[1][None], pts[None])
    o_out, o_st = olk.lk(fr[0], fr[1], pts)
    assert st[0][0] == 0 and np.array_equal(st[0], o_st)
    assert np.array_equal(out[0][o_st == 1], o_out[o_st == 1])
    f.close()


def check_state(gs, i, os_):
    uv = gs["update_vec"][i]
    assert list(uv) == list(os_["update_vec"])
    live = uv == 1
    assert np.array_equal(gs["feat_cur"][i][live], os_["feat_cur"][live])
    assert np.array_equal(gs["prev_corners"][i][live], os_["prev_corners"][live])


@pytest.mark.parametrize("seq", GOLD["seq"], ids=lambda s: "%dx%d_r%d_n%d" % (s["w"], s["h"], s["rate"], s["nf"]))
def test_calcflow_sequence_single_stream(pkg, seq):
    """aof_lk_calc_flow == the oracle's OpticalFlowOpenCV::calcFlow, call by call (state included)."""
    T = len(seq["rows"])
    fr = pkg.synth.lk_sequence(T, seq["w"], seq["h"], seq["dx"], seq["dy"], seed=seq["seed"])
    f = pkg.LkFlow(FX, FY, seq["rate"], seq["w"], seq["h"], seq["nf"])
    o = olk.Tracker(FX, FY, seq["rate"], seq["w"], seq["h"], seq["nf"])
    for k, row in enumerate(seq["rows"]):
        t = 1000 + 13333 * k
        q, dt, fx, fy = f.calc_flow(fr[k], t)
        oq, odt, ofx, ofy = o.calc_flow(fr[k], t)
        assert (q, fx, fy) == (oq, ofx, ofy), k
        if oq >= 0 or seq["rate"] <= 0:
            assert dt == odt
        else:
            assert dt == -123456                     # untouched when not publishing
        check_state(f.state(), 0, o.state())
        assert q == row["q"] and abs(fx - row["fx"]) <= ANG_TOL and abs(fy - row["fy"]) <= ANG_TOL   # vs cv2 restatement
    f.close()


def test_batch_streams_equal_independent_oracle_runs(pkg):
    """N streams in one handle == N single-stream oracle objects; batch order does not matter."""
    w, h, n, T, nf = 128, 128, 24, 7, 20
    rng = np.random.default_rng(17)
    vel = rng.integers(-4, 5, (n, 2))
    seqs = [pkg.synth.lk_sequence(T, w, h, int(vel[i, 0]), int(vel[i, 1]), seed=100 + i) for i in range(n)]
    seqs[3] = [np.full((h, w), 40, np.uint8)] * T                                     # a stream that never finds corners
    frames = np.stack([np.stack([seqs[i][k] for i in range(n)]) for k in range(T)])   # [T, n, h, w]
    tstep = rng.integers(9000, 30000, n).astype(np.uint32)
    perm = rng.permutation(n)
    for rate in (-1, 15):
        f = pkg.LkFlow(FX, FY, rate, w, h, nf, n_streams=n)
        g = pkg.LkFlow(FX, FY, rate, w, h, nf, n_streams=n)
        orc = [olk.Tracker(FX, FY, rate, w, h, nf) for _ in range(n)]
        for k in range(T):
            t_us = (1000 + tstep * k).astype(np.uint32)
            out, raw = f.calc_flow_batch(frames[k], t_us, want_raw=True)
            outp = g.calc_flow_batch(frames[k][perm], t_us[perm])
            assert np.array_equal(out[perm], outp)
            gs = f.state()
            for i in range(n):
                q, dt, fx, fy = orc[i].calc_flow(frames[k, i], int(t_us[i]))
                os_ = orc[i].state()
                assert out["quality"][i] == q and out["flow_x"][i] == np.float32(fx) and out["flow_y"][i] == np.float32(fy), (k, i)
                if q >= 0 or rate <= 0:
                    assert out["dt_us"][i] == dt
                assert (raw["inliers"][i], raw["fast_bins"][i], raw["active"][i]) == (os_["inliers"], os_["fast_bins"], os_["active"])
                assert raw["px_x"][i] == np.float32(os_["px_x"]) and raw["px_y"][i] == np.float32(os_["px_y"])
                check_state(gs, i, os_)
        f.close(); g.close()


def test_translation_recovered_640x480(pkg):
    """BASELINE config 2 size: the flow of a translating texture is the translation (1e-3 px)."""
    n, T = 4, 5
    vel = [(-3, 2), (4, 1), (0, -5), (2, 2)]
    seqs = [pkg.synth.lk_sequence(T, 640, 480, vx, vy, seed=200 + i) for i, (vx, vy) in enumerate(vel)]
    f = pkg.LkFlow(FX, FY, -1, 640, 480, 20, n_streams=n)
    for k in range(T):
        out, raw = f.calc_flow_batch(np.stack([s[k] for s in seqs]), np.full(n, 1000 + 10000 * k, np.uint32), want_raw=True)
        if k == 0:
            assert (out["quality"] == 0).all() and (raw["fast_bins"] == 16).all()
            continue
        assert (out["quality"] > 100).all()
        for i, (vx, vy) in enumerate(vel):
            assert abs(raw["px_x"][i] - vx) < 1e-3 and abs(raw["px_y"][i] - vy) < 1e-3
            assert abs(out["flow_x"][i] - np.arctan2(vx, FX)) < ANG_TOL and abs(out["flow_y"][i] - np.arctan2(vy, FY)) < ANG_TOL
    f.close()


def test_reset_and_errors(pkg):
    f = pkg.LkFlow(FX, FY, -1, 64, 64)
    fr = pkg.synth.lk_sequence(3, 64, 64, 1, 1, seed=1)
    a = [f.calc_flow(x, 1000 * (k + 1)) for k, x in enumerate(fr)]
    f.reset()
    b = [f.calc_flow(x, 1000 * (k + 1)) for k, x in enumerate(fr)]
    assert a == b
    f.close()
    with pytest.raises(pkg.AofError):
        pkg.LkFlow(FX, FY, -1, 64, 64, num_features=101)
    with pytest.raises(pkg.AofError):
        pkg.LkFlow(FX, FY, -1, 8, 8)


def test_undistort_branch_matches_oracle(pkg):
    """setCameraMatrix + setCameraDistortion switch on cv::undistortPoints (flow_opencv.cpp:108-119)."""
    w = h = 128
    fr = pkg.synth.lk_sequence(7, w, h, -2, 3, seed=77)
    f = pkg.LkFlow(FX, FY, -1, w, h, 20)
    o = olk.Tracker(FX, FY, -1, w, h, 20)
    f.set_camera_matrix(FX, FY, 64.0, 64.0)
    o.set_camera_matrix(FX, FY, 64.0, 64.0)
    r0 = f.calc_flow(fr[0], 1000)                    # only one setter called: still the plain path
    assert r0 == o.calc_flow(fr[0], 1000)
    f.set_camera_distortion(-0.2, 0.05, 0.01, 0.001, -0.002)
    o.set_camera_distortion(-0.2, 0.05, 0.01, 0.001, -0.002)
    for k in range(1, 7):
        assert f.calc_flow(fr[k], 1000 + 9000 * k) == o.calc_flow(fr[k], 1000 + 9000 * k), k
        check_state(f.state(), 0, o.state())
    f.close()


@pytest.mark.parametrize("w,h,nf", [(1280, 720, 40), (1024, 768, 100)])
def test_large_frames_match_oracle(pkg, w, h, nf):
    """Bins of 320 x 180 / 256 x 192 px hold more FAST survivors than the selection kernel stages in shared memory
    (the tail is read from global memory), the banded FAST kernel runs its run-time-pitch form, 100 = the slot limit."""
    n, T = 2, 4
    seqs = [pkg.synth.lk_sequence(T, w, h, 3 - 5 * i, -2 + 3 * i, seed=700 + i) for i in range(n)]
    f = pkg.LkFlow(FX, FY, -1, w, h, nf, n_streams=n)
    orc = [olk.Tracker(FX, FY, -1, w, h, nf) for _ in range(n)]
    for k in range(T):
        t_us = np.full(n, 1000 + 20000 * k, np.uint32)
        out, raw = f.calc_flow_batch(np.stack([s[k] for s in seqs]), t_us, want_raw=True)
        gs = f.state()
        for i in range(n):
            q, dt, fx, fy = orc[i].calc_flow(seqs[i][k], int(t_us[i]))
            os_ = orc[i].state()
            assert out["quality"][i] == q and out["flow_x"][i] == np.float32(fx) and out["flow_y"][i] == np.float32(fy), (k, i)
            assert (raw["inliers"][i], raw["fast_bins"][i], raw["active"][i]) == (os_["inliers"], os_["fast_bins"], os_["active"])
            check_state(gs, i, os_)
    f.close()


def _run_sequences(pkg, frames, w, h, nf, rate=-1):
    n, T = frames.shape[1], frames.shape[0]
    f = pkg.LkFlow(FX, FY, rate, w, h, nf, n_streams=n)
    outs = []
    for k in range(T):
        res, raw = f.calc_flow_batch(frames[k], np.full(n, 1000 + 13333 * k, np.uint32), want_raw=True)
        outs.append((res.copy(), raw.copy(), {key: v.copy() for key, v in f.state().items()}))
    stats = f.lazy_stats()
    f.close()
    return outs, stats


@pytest.mark.parametrize("mode", ["2", "1"])
@pytest.mark.parametrize("w,h,nf", [(640, 480, 20), (320, 240, 50), (128, 128, 20), (256, 192, 100)])
def test_lazy_fast_equals_complete_path(pkg, w, h, nf, mode, monkeypatch):
    """The two shortcuts of the detector -- AOF_FAST_LAZY=2 (default): lk_fast_band_kernel scoring only the pixels outside
    the exclusion squares of the active points; =1: lk_fast_pick_kernel (bound, pool, exact scores of the candidates
    only) -- against the complete keypoint lists (AOF_FAST_LAZY=0), byte for byte: outputs, diagnostics and the whole tracker state, frame by frame; includes
    a scene cut (every feature lost: all 16 bins requested again) and flat frames (no corner anywhere)."""
    n, T = 12, 9
    rng = np.random.default_rng(w + nf)
    seqs = [pkg.synth.lk_sequence(T, w, h, int(rng.integers(-4, 5)), int(rng.integers(-4, 5)), seed=900 + i,
                                  sigma=float(rng.choice([1.2, 2.0, 3.5]))) for i in range(n)]
    frames = np.stack([np.stack([seqs[i][k] for i in range(n)]) for k in range(T)])
    frames[5, :4] = frames[0, 4:8]                       # scene cut for four streams
    frames[6, 8] = 90                                    # a flat frame
    frames[:, 9] = (frames[:, 9] // 8) + 100             # a low-contrast stream: bounds below any learnt floor
    monkeypatch.setenv("AOF_FAST_LAZY", mode)
    lazy, stats = _run_sequences(pkg, frames, w, h, nf)
    monkeypatch.setenv("AOF_FAST_LAZY", "0")
    full, stats0 = _run_sequences(pkg, frames, w, h, nf)
    assert stats0 == (0, 0, 0, 0) and (stats[0] > 0 or mode == "2"), (stats, stats0)
    for k in range(T):
        assert lazy[k][0].tobytes() == full[k][0].tobytes(), k
        assert lazy[k][1].tobytes() == full[k][1].tobytes(), k
        for key in full[k][2]:
            assert np.array_equal(lazy[k][2][key], full[k][2][key]), (k, key)
    # and against the oracle for two of the streams
    for i in (0, 9):
        o = olk.Tracker(FX, FY, -1, w, h, nf)
        for k in range(T):
            q, dt, fx, fy = o.calc_flow(frames[k, i], 1000 + 13333 * k)
            r = lazy[k][0]
            assert (r["quality"][i], r["flow_x"][i], r["flow_y"][i]) == (q, np.float32(fx), np.float32(fy)), (i, k)


@pytest.mark.parametrize("mode", ["2", "1"])
def test_lazy_fast_second_chance_goes_through_the_complete_path(pkg, mode, monkeypatch):
    """sparse blobs: fewer acceptable keypoints than requests, so initMorePoints' second-chance pass runs
    (trackFeatures.cpp:266-304); the lazy selection must hand those streams to the complete path and still equal the oracle"""
    from test_ref_lk import blob_sequence
    w, h, nf, T = 320, 240, 33, 6
    seqs = [blob_sequence(T, w, h, 2, -1, 41 + i) for i in range(4)]
    frames = np.stack([np.stack([seqs[i][k] for i in range(4)]) for k in range(T)])
    monkeypatch.setenv("AOF_FAST_LAZY", mode)
    got, stats = _run_sequences(pkg, frames, w, h, nf)
    assert stats[3] > 0, stats                           # the fallback really ran
    for i in range(4):
        o = olk.Tracker(FX, FY, -1, w, h, nf)
        for k in range(T):
            q, dt, fx, fy = o.calc_flow(frames[k, i], 1000 + 13333 * k)
            r = got[k][0]
            assert (r["quality"][i], r["flow_x"][i], r["flow_y"][i]) == (q, np.float32(fx), np.float32(fy)), (i, k)
            check_state(got[k][2], i, o.state())
