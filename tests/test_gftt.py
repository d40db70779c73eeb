"""Shi-Tomasi / goodFeaturesToTrack (the selector north_star names).  Oracle = oracle/lk_oracle.c
(olk_min_eig, olk_gftt), pinned to cv2 4.13: the response within 5e-8 absolute (cv2's vectorised column
filter may fuse multiply-adds), the selected corner sets identical.  GPU vs oracle: bit-exact."""
import numpy as np
import pytest

from oracle import lk as olk

try:
    import cv2
    cv2.setNumThreads(1)
except Exception:  # pragma: no cover
    cv2 = None

CASES = [(640, 480, 0), (128, 128, 1), (75, 61, 2)]
PARAMS = [(20, 0.01, 10.0), (100, 0.05, 5.0), (50, 0.01, 0.0)]


@pytest.mark.skipif(cv2 is None, reason="cv2 not importable")
@pytest.mark.parametrize("w,h,seed", CASES)
def test_oracle_vs_cv2(pkg, w, h, seed):
    img = pkg.synth.lk_sequence(1, w, h, 1, 1, seed=seed)[0]
    eig = olk.min_eig(img)
    ref = cv2.cornerMinEigenVal(img, 3, ksize=3)
    assert np.abs(eig - ref).max() <= 5e-8
    for n, q, md in PARAMS:
        c = cv2.goodFeaturesToTrack(img, n, q, md)
        c = c.reshape(-1, 2) if c is not None else np.zeros((0, 2), np.float32)
        assert np.array_equal(olk.gftt(img, n, q, md), c)


def test_oracle_edge_cases():
    flat = np.full((40, 50), 9, np.uint8)
    assert len(olk.gftt(flat, 10, 0.01, 3.0)) == 0 and (olk.min_eig(flat) == 0).all()
    one = flat.copy(); one[20:30, 20:30] = 200          # a bright square: its four corners are the strongest responses
    c = olk.gftt(one, 4, 0.5, 5.0)
    assert len(c) == 4 and set(map(tuple, np.round((c - 24.5) / 5).astype(int).tolist())) == {(-1, -1), (-1, 1), (1, -1), (1, 1)}


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,seed", CASES + [(322, 242, 3)])
def test_gpu_vs_oracle(pkg, w, h, seed):
    imgs = np.stack([pkg.synth.lk_sequence(1, w, h, 1, 1, seed=seed + 10 * k)[0] for k in range(3)])
    imgs[2, : h // 2] = 60                                # half flat
    g = pkg.Gftt(w, h)
    eig = g.min_eig(imgs)
    for i in range(len(imgs)):
        assert np.array_equal(eig[i], olk.min_eig(imgs[i])), i
    for n, q, md in PARAMS:
        got = g.detect(imgs, n, q, md)
        for i in range(len(imgs)):
            assert np.array_equal(got[i], olk.gftt(imgs[i], n, q, md)), (i, n, q, md)
    g.close()
