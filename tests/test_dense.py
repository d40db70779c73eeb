"""Dense per-block SAD field (BASELINE config 4).  The oracle (oracle/sad_oracle.c: orc_dense_field) is the
reference's three per-block primitives with 32-bit offsets; it is pinned to the unmodified reference objects on
tiles the reference can address (<= 255 px pitch), then the CUDA kernel is compared bit-exactly with it."""
import numpy as np
import pytest

from oracle import sad as osad


def translated_pair(pkg, w, h, dx, dy, seed, half=False):
    rng = np.random.default_rng(seed)
    m = 12
    tex = pkg.synth.box_blur3(rng.integers(0, 256, (h + 2 * m, w + 2 * m), dtype=np.uint8))
    prev = np.ascontiguousarray(tex[m:m + h, m:m + w])
    cur = tex[m - dy:m - dy + h, m - dx:m - dx + w]
    if half:
        cur = ((cur.astype(np.uint16) + tex[m - dy:m - dy + h, m - dx - 1:m - dx - 1 + w] + 1) >> 1).astype(np.uint8)
    return prev, np.ascontiguousarray(cur)


def test_dense_oracle_vs_reference_primitives(pkg, port, ref):
    """every block of the oracle's field == the reference's own compute_diff / compute_sad_8x8 / compute_subpixel"""
    w, h, s = 96, 80, 4
    prev, cur = translated_pair(pkg, w, h, 3, -2, 5)
    cur[40:, :] = np.random.default_rng(1).integers(0, 256, (h - 40, w), dtype=np.uint8)   # lower half: uncorrelated
    prev[:, 60:] = 128                                                                       # right part: no texture
    f = osad.dense_field(port, prev, cur, s, 30, 3000)
    i_lo = (s + 1 + 7) // 8
    assert f.shape == ((h - s - 9) // 8 - i_lo + 1, (w - s - 9) // 8 - i_lo + 1)
    for bj in range(f.shape[0]):
        for bi in range(f.shape[1]):
            x, y, o = 8 * (i_lo + bi), 8 * (i_lo + bj), f[bj, bi]
            assert o["texture"] == ref.texture(prev, x, y)
            if o["texture"] < 30:
                assert o["flags"] == 0 and o["best_sad"] == 0xFFFFFFFF
                continue
            sads = {(ii, jj): ref.sad8x8(prev, cur, x, y, x + ii, y + jj) for jj in range(-s, s + 1) for ii in range(-s, s + 1)}
            best = min(sads.values())
            first = next(k for k in sorted(sads, key=lambda k: (k[1], k[0])) if sads[k] == best)   # jj outer, ii inner
            assert (o["best_sad"], o["dx"], o["dy"]) == (best, first[0], first[1])
            if best < 3000:
                acc = ref.subpixel(prev, cur, x, y, x + first[0], y + first[1])
                sub, m = 8, best
                for k in range(8):
                    if acc[k] < m:
                        m, sub = acc[k], k
                assert o["flags"] == 3 and o["subdir"] == sub
            else:
                assert o["flags"] == 1 and o["subdir"] == 8
    assert (f["flags"] == 3).any() and (f["flags"] == 1).any() and (f["flags"] == 0).any()


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,s", [(1280, 720, 8), (640, 480, 8), (320, 240, 6), (96, 80, 4), (132, 100, 8)])
def test_dense_gpu_vs_oracle(pkg, port, w, h, s):
    pairs = [translated_pair(pkg, w, h, s, -s, 11), translated_pair(pkg, w, h, -3, 2, 12, half=True),
             translated_pair(pkg, w, h, 0, 0, 13)]
    pairs[2][1][h // 2:, :] = np.random.default_rng(2).integers(0, 256, (h - h // 2, w), dtype=np.uint8)
    pairs[2][0][:, : w // 3] = 77
    prev = np.stack([p[0] for p in pairs]); cur = np.stack([p[1] for p in pairs])
    d = pkg.SadDense(w, h, s, 30, 3000)
    got = d.compute(prev, cur)
    for i in range(len(pairs)):
        want = osad.dense_field(port, prev[i], cur[i], s, 30, 3000)
        assert got[i].shape == want.shape
        for name in ("texture", "flags", "best_sad", "subdir"):
            assert np.array_equal(got[i][name], want[name]), (i, name)
        live = (want["flags"] & 1) == 1
        assert np.array_equal(got[i]["dx"][live], want["dx"][live]) and np.array_equal(got[i]["dy"][live], want["dy"][live])
    if (w, h, s) == (1280, 720, 8):
        assert (d.nbx, d.nby, d.x0, d.y0) == (156, 86, 16, 16)
        acc = got[0]["flags"] == 3
        assert acc.mean() > 0.99 and (got[0]["dx"][acc] == 8).all() and (got[0]["dy"][acc] == -8).all()
        acc = got[1]["flags"] == 3
        assert np.isin(got[1]["dx"][acc], (-3, -2)).all() and (got[1]["dy"][acc] == 2).all() and (got[1]["subdir"][acc] != 8).mean() > 0.9
    d.close()
