"""CPU: pin the SAD oracle (oracle/sad_oracle.c) against the committed golden vectors
(outputs of the unmodified reference, tests/golden/sad_golden.json) and, where
oracle/_ref is present, against the reference objects directly."""
import json
import os

import numpy as np
import pytest

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "sad_golden.json")))


@pytest.fixture(scope="module")
def synth():
    import __graft_entry__ as entry
    return entry.load_package().synth


@pytest.fixture(scope="module")
def lcg(synth):
    tex = synth.lcg_texture()
    return tex, synth.crop(tex, 100, 100), synth.crop(tex, 98, 101)


def test_kat_primitives(port, lcg):
    _, A, B = lcg
    k = GOLD["kat"]
    assert port.texture(A, 5, 5) == k["diff_A_5_5"] == 340
    assert port.sad8x8(A, B, 5, 5, 5, 5) == k["sad_A_B_5_5_5_5"] == 1568
    assert port.sad8x8(A, B, 5, 5, 7, 4) == k["sad_A_B_5_5_7_4"] == 0
    assert list(port.subpixel(A, B, 5, 5, 7, 4)) == k["subpixel_A_B_5_5_7_4"] == [549, 681, 411, 673, 527, 617, 427, 656]


def test_kat_shifts(port, synth, lcg):
    tex, A, B = lcg
    for e in GOLD["kat"]["lcg_shifts"]:
        Bs = synth.crop(tex, 100 - e["dx"], 100 - e["dy"])
        fx, fy, q = port.compute_flow(A, Bs, e["s"])
        assert (fx, fy, q) == (e["fx"], e["fy"], e["q"])
        if abs(e["dx"]) <= e["s"] and abs(e["dy"]) <= e["s"]:
            assert (fx, fy, q) == (e["dx"], e["dy"], 255)
    Bh = ((B.astype(np.int32) + synth.crop(tex, 97, 101) + 1) // 2).astype(np.uint8)
    assert list(port.compute_flow(A, Bh, 4)) == GOLD["kat"]["half_pixel"] == [2.5, -1.0, 255]
    rng = np.random.default_rng(99)
    low = (128 + rng.integers(0, 3, (64, 64))).astype(np.uint8)
    assert list(port.compute_flow(low, np.roll(low, 1, axis=1), 4)) == GOLD["kat"]["low_texture"] == [0.0, 0.0, 0]


@pytest.mark.parametrize("case", GOLD["pairs"], ids=lambda c: "seed%d_s%d" % (c["seed"], c["search"]))
def test_bulk_pairs_vs_golden(port, synth, case):
    prev, cur, _ = synth.sad_pairs(case["n"], max_shift=case["max_shift"], seed=case["seed"],
                                   half_pixel_every=case["half_pixel_every"])
    fx, fy, q = port.compute_flow_batch(prev, cur, case["search"], 30, case["val_thr"])
    assert np.array_equal(fx, np.array(case["fx"], np.float32))
    assert np.array_equal(fy, np.array(case["fy"], np.float32))
    assert np.array_equal(q, np.array(case["q"], np.int32))


@pytest.mark.parametrize("seq", GOLD["seq"], ids=lambda s: "rate%d" % s["rate"])
def test_calcflow_sequences_vs_golden(port, synth, lcg, seq):
    tex = lcg[0]
    o = port.px4(216.6677, 216.2457, seq["rate"], 64, 64, 4, 30, 3000)
    for t, row in enumerate(seq["rows"]):
        img = synth.crop(tex, 100 - 2 * t, 100 + t)
        q, dt, fx, fy, *_ = o.calc_flow(img, 1000 + seq["dt_step"] * t)
        assert q == row["q"]
        if row["fx"] is None:
            assert np.isnan(fx) and np.isnan(fy) and dt == -123456   # outputs untouched on the priming call
        else:
            assert fx == np.float32(row["fx"]) and fy == np.float32(row["fy"])
        assert dt == row["dt"]


def test_port_vs_reference_random(port, ref, synth):
    """Direct cross-check against the reference objects: translated textures, raw noise, odd thresholds."""
    rng = np.random.default_rng(5)
    for s in (4, 6, 8):
        prev, cur, _ = synth.sad_pairs(60, max_shift=s + 1, seed=100 + s, half_pixel_every=2)
        for i in range(60):
            assert port.compute_flow(prev[i], cur[i], s) == ref.compute_flow(prev[i], cur[i], s)
    for _ in range(60):
        a = rng.integers(0, 256, (64, 64), dtype=np.uint8)
        b = rng.integers(0, 256, (64, 64), dtype=np.uint8)
        vt = int(rng.integers(3000, 9000)); ft = int(rng.integers(0, 4000))
        assert port.compute_flow(a, b, 4, ft, vt) == ref.compute_flow(a, b, 4, ft, vt)
        x, y = int(rng.integers(5, 45)), int(rng.integers(5, 45))
        assert port.texture(a, x, y) == ref.texture(a, x, y)
        assert port.sad8x8(a, b, x, y, x + 3, y - 2) == ref.sad8x8(a, b, x, y, x + 3, y - 2)
        assert list(port.subpixel(a, b, x, y, x - 1, y + 2)) == list(ref.subpixel(a, b, x, y, x - 1, y + 2))


def test_port_vs_reference_other_widths(port, ref):
    rng = np.random.default_rng(6)
    for w in (32, 48, 96, 128, 252):
        a = rng.integers(0, 256, (w, w), dtype=np.uint8)
        b = np.roll(np.roll(a, 2, axis=1), -1, axis=0)
        for s in (4, 6):
            if w < 2 * (s + 1) + 9:
                continue
            assert port.compute_flow(a, b, s, 30, 6000) == ref.compute_flow(a, b, s, 30, 6000)


def test_px4_setters_keep_state_port_vs_reference(pkg, port, ref):
    """the reference's setters only assign (include/optical_flow.hpp:76-78): the port's set_params must follow the
    reference objects through a change of focal length and output rate in mid-sequence"""
    tex = pkg.synth.lcg_texture()
    frames = [pkg.synth.crop(tex, 100 - 2 * t, 100 + t) for t in range(14)]
    a = port.px4(216.6677, 216.2457, 15, 64, 64, 4, 30, 3000)
    b = ref.px4(216.6677, 216.2457, 15, 64, 64, 4, 30, 3000)
    retune = {4: (300.0, 310.0, 15), 8: (300.0, 310.0, -1), 11: (216.0, 216.0, 20)}
    for k, f in enumerate(frames):
        if k in retune:
            a.set_params(*retune[k]); b.set_params(*retune[k])
        ra = a.calc_flow(f, 1000 + 13333 * k)[:4]
        rb = b.calc_flow(f, 1000 + 13333 * k)
        assert ra[0] == rb[0] and ra[1] == rb[1], k
        if k:
            assert ra[2:] == rb[2:], k
