"""Pins oracle/lk_oracle.c (the CPU restatement of the OpenCV path) against
  * tests/golden/lk_golden.json -- outputs of cv2 4.13 and of the cv2-driven literal restatement
    of trackFeatures.cpp + flow_opencv.cpp (tests/golden/make_lk_golden.py), and
  * cv2 itself when it is importable (it is in this image).
Bit-exact for pyramid levels, FAST scores and LK status; LK positions within 2e-4 px of cv2
(cv2 adds f32 terms in SIMD-lane order; the oracle sums exactly)."""
import hashlib
import json
import os

import numpy as np
import pytest

from oracle import lk

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "lk_golden.json")))
FX, FY = 216.6677, 216.2457
PX_TOL = 2e-4                      # px, oracle vs cv2 positions
ANG_TOL = 1e-3 / 216.0             # rad == 1e-3 px of flow (north_star tolerance)

try:
    import cv2
    cv2.setNumThreads(1)
except Exception:                  # pragma: no cover
    cv2 = None


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


@pytest.mark.parametrize("case", GOLD["pyr"], ids=lambda c: "%dx%d" % (c["w"], c["h"]))
def test_pyr_down_golden(pkg, case):
    a = pkg.synth.lk_sequence(1, case["w"], case["h"], 1, 1, seed=case["seed"])[0]
    for want in case["levels"]:
        a = lk.pyr_down(a)
        assert sha(a) == want


def test_pyr_levels_rule():
    assert lk.pyr_top_level(64, 64) == 2 and lk.pyr_top_level(128, 128) == 3 and lk.pyr_top_level(640, 480) == 3
    assert lk.pyr_top_level(20, 20) == 1 and lk.pyr_top_level(19, 40) == 1 and lk.pyr_top_level(18, 40) == 0


@pytest.mark.parametrize("case", GOLD["lk"], ids=lambda c: "%dx%d" % (c["w"], c["h"]))
def test_lk_golden(pkg, case):
    fr = pkg.synth.lk_sequence(2, case["w"], case["h"], case["dx"], case["dy"], seed=case["seed"])
    out, st = lk.lk(fr[0], fr[1], np.array(case["pts"], np.float32))
    want_st = np.array(case["status"], np.uint8)
    assert np.array_equal(st, want_st)
    d = np.abs(out - np.array(case["next"]))[want_st == 1]
    assert d.max() <= PX_TOL


@pytest.mark.parametrize("case", GOLD["fast"], ids=lambda c: "roi%d_%d" % (c["roi"][0], c["roi"][1]))
def test_fast_golden(pkg, case):
    img = pkg.synth.lk_sequence(1, 640, 480, 1, 1, seed=0)[0]
    y0, x0, h, w = case["roi"]
    sc = lk.fast(img[y0:y0 + h, x0:x0 + w])
    assert int((sc > 0).sum()) == case["n"] and sha(sc) == case["sha"]


@pytest.mark.parametrize("seq", GOLD["seq"], ids=lambda s: "%dx%d_r%d_n%d" % (s["w"], s["h"], s["rate"], s["nf"]))
def test_calcflow_sequence_golden(pkg, seq):
    fr = pkg.synth.lk_sequence(len(seq["rows"]), seq["w"], seq["h"], seq["dx"], seq["dy"], seed=seq["seed"])
    t = lk.Tracker(FX, FY, seq["rate"], seq["w"], seq["h"], seq["nf"])
    for k, row in enumerate(seq["rows"]):
        q, dt, fx, fy = t.calc_flow(fr[k], 1000 + 13333 * k)
        s = t.state()
        assert q == row["q"] and s["inliers"] == row["inliers"] and s["fast_bins"] == row["fast_bins"]
        assert list(s["update_vec"]) == row["update"]
        if row["dt"] is not None:
            assert dt == row["dt"]
        assert abs(fx - row["fx"]) <= ANG_TOL and abs(fy - row["fy"]) <= ANG_TOL
        assert abs(s["px_x"] - row["px"][0]) <= 1e-3 and abs(s["px_y"] - row["px"][1]) <= 1e-3


def test_first_frame_and_static_scene(pkg):
    fr = pkg.synth.lk_sequence(1, 128, 128, 0, 0, seed=9)[0]
    t = lk.Tracker(FX, FY, -1, 128, 128, 20)
    q, dt, fx, fy = t.calc_flow(fr, 5000)
    assert (q, dt, fx, fy) == (0, 5000, 0.0, 0.0) and t.state()["fast_bins"] == 16
    q, dt, fx, fy = t.calc_flow(fr, 15000)             # identical frame: zero flow, every feature an inlier or sigma == 0
    assert dt == 10000 and abs(fx) < 1e-6 and abs(fy) < 1e-6
    flat = np.full((128, 128), 77, np.uint8)            # no corners anywhere: all requests fail, quality 0
    t2 = lk.Tracker(FX, FY, -1, 128, 128, 20)
    for k in range(3):
        assert t2.calc_flow(flat, 1000 * k)[0] == 0
    assert (t2.state()["update_vec"] == 2).all()


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_primitives_vs_cv2_live(pkg):
    fr = pkg.synth.lk_sequence(2, 320, 240, 2, -3, seed=12)
    a = fr[0]
    for _ in range(3):
        assert np.array_equal(lk.pyr_down(a), cv2.pyrDown(a))
        a = lk.pyr_down(a)
    d = lk.scharr(fr[0])
    assert np.array_equal(d[..., 0], cv2.Scharr(fr[0], cv2.CV_16S, 1, 0))
    assert np.array_equal(d[..., 1], cv2.Scharr(fr[0], cv2.CV_16S, 0, 1))
    rng = np.random.default_rng(5)
    pts = rng.uniform([-5, -5], [325, 245], (300, 2)).astype(np.float32)
    o, s = lk.lk(fr[0], fr[1], pts)
    c, cs, _ = cv2.calcOpticalFlowPyrLK(fr[0], fr[1], pts.reshape(-1, 1, 2), None, winSize=(9, 9), maxLevel=3)
    assert np.array_equal(s, cs.reshape(-1))
    assert np.abs(o - c.reshape(-1, 2))[s == 1].max() <= PX_TOL
    det = cv2.FastFeatureDetector_create(10, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    roi = fr[0][60:120, 80:160]
    m = np.zeros(roi.shape, np.uint8)
    for k in det.detect(np.ascontiguousarray(roi)):
        m[int(k.pt[1]), int(k.pt[0])] = int(k.response)
    assert np.array_equal(lk.fast(roi), m)


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_calcflow_vs_cv2_restatement_live(pkg):
    from cv2_restatement import Cv2OpticalFlowOpenCV
    for (w, h, dx, dy, seed, rate, nf) in [(160, 120, 1, 2, 21, -1, 20), (128, 128, -2, 1, 22, 15, 50), (96, 96, 3, 3, 23, -1, 7)]:
        fr = pkg.synth.lk_sequence(8, w, h, dx, dy, seed=seed)
        A = lk.Tracker(FX, FY, rate, w, h, nf)
        B = Cv2OpticalFlowOpenCV(FX, FY, rate, w, h, nf)
        for k in range(8):
            ra = A.calc_flow(fr[k], 1000 + 13333 * k)
            rb = B.calc_flow(fr[k], 1000 + 13333 * k)
            assert ra[0] == rb[0] and list(A.state()["update_vec"]) == B.update
            assert abs(ra[2] - rb[2]) <= ANG_TOL and abs(ra[3] - rb[3]) <= ANG_TOL


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_undistort_vs_cv2_live():
    rng = np.random.default_rng(0)
    for cam, dc in [((216.6677, 216.2457, 64, 64), (-0.2, 0.05, 0.001, -0.002, 0.01)),
                    ((500.5, 498.1, 320.2, 239.7), (0.12, -0.3, 0.0, 0.0, 0.2)), ((216, 216, 64, 64), (0, 0, 0, 0, 0))]:
        pts = np.concatenate([rng.uniform(-20, 700, (300, 2)), [[-100, -100]]]).astype(np.float32)
        K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
        u = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, np.array([dc], np.float32)).reshape(-1, 2)
        want = np.stack([u[:, 0] * np.float32(cam[0]) + np.float32(cam[2]), u[:, 1] * np.float32(cam[1]) + np.float32(cam[3])], 1)
        assert np.array_equal(olk_undistort(cam, dc, pts), want.astype(np.float32))


def olk_undistort(cam, dc, pts):
    return lk.undistort(cam, dc, pts)


def test_undistort_known_answers():
    # identity distortion is the identity map up to the float round trip; golden values for a distorted camera
    cam, dc = (216.6677, 216.2457, 64.0, 64.0), (-0.2, 0.05, 0.001, -0.002, 0.01)
    pts = np.array([[64, 64], [0, 0], [127, 127], [-100, -100], [100.5, 20.25]], np.float32)
    assert np.abs(lk.undistort(cam, (0, 0, 0, 0, 0), pts) - pts).max() < 1e-4
    got = lk.undistort(cam, dc, pts)
    assert np.array_equal(got[0], [64, 64]) and (np.abs(got[1:] - pts[1:]).max(1) > 0.01).all()


def test_lk_final_position_retest_known_answer(pkg):
    """calcOpticalFlowPyrLK re-tests the window origin of the FINAL position at level 0 when the error vector is
    requested (trackFeatures.cpp:94,98 does; cv2 always does): a point whose last update crosses the range ends with
    status 0.  Found by searching the band x = cols + 4 - dx; cv2 4.13 answers 0 for it, a tracker without the re-test 1."""
    fr = pkg.synth.lk_sequence(2, 160, 120, 3, 0, seed=75)
    pts = np.array([[160.72454833984375, 53.756690979003906], [80.0, 60.0]], np.float32)
    out, st = lk.lk(fr[0], fr[1], pts)
    assert list(st) == [0, 1]
    ox = np.floor(out[0, 0] - 4)
    assert ox < -9 or ox >= 160                     # it was the final update that left the range
    if cv2 is not None:
        c, cs, _ = cv2.calcOpticalFlowPyrLK(fr[0], fr[1], pts.reshape(-1, 1, 2), None, winSize=(9, 9), maxLevel=3)
        assert list(cs.reshape(-1)) == [0, 1] and np.abs(c.reshape(-1, 2) - out).max() <= PX_TOL


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
def test_lk_status_at_the_range_boundary_vs_cv2_live(pkg):
    rng = np.random.default_rng(3)
    for seed, (dx, dy) in enumerate([(3, 0), (-3, 0), (0, 3), (0, -3)]):
        fr = pkg.synth.lk_sequence(2, 160, 120, dx, dy, seed=75 + seed)
        n = 1500
        band = rng.uniform(-0.3, 0.3, n)
        if dx:
            x0 = 160 + 4 - dx if dx > 0 else -9 + 4 - dx
            pts = np.stack([x0 + band, rng.uniform(10, 110, n)], 1)
        else:
            y0 = 120 + 4 - dy if dy > 0 else -9 + 4 - dy
            pts = np.stack([rng.uniform(10, 150, n), y0 + band], 1)
        pts = pts.astype(np.float32)
        o, s = lk.lk(fr[0], fr[1], pts)
        c, cs, _ = cv2.calcOpticalFlowPyrLK(fr[0], fr[1], pts.reshape(-1, 1, 2), None, winSize=(9, 9), maxLevel=3)
        assert np.array_equal(s, cs.reshape(-1))
