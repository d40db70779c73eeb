"""Generates tests/golden/sad_golden.json by EXECUTING THE UNMODIFIED REFERENCE
(oracle/_ref/libaof_ref.so, built from /root/reference by oracle/Makefile) on the
deterministic inputs of aero-optical-flow_b200/synth.py.  Run in the authoring
container only:  python tests/golden/make_sad_golden.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402
from oracle import sad  # noqa: E402

synth = entry.load_package().synth
R = sad.Ref()
G = {"source": "reference objects @23dad58 / PX4 OpticalFlow 39732aa, g++ -O3", "kat": {}, "pairs": [], "seq": []}

tex = synth.lcg_texture()
A = synth.crop(tex, 100, 100)
B = synth.crop(tex, 98, 101)
G["kat"]["diff_A_5_5"] = int(R.texture(A, 5, 5))
G["kat"]["sad_A_B_5_5_5_5"] = int(R.sad8x8(A, B, 5, 5, 5, 5))
G["kat"]["sad_A_B_5_5_7_4"] = int(R.sad8x8(A, B, 5, 5, 7, 4))
G["kat"]["subpixel_A_B_5_5_7_4"] = [int(v) for v in R.subpixel(A, B, 5, 5, 7, 4)]
shifts = []
for s in (4, 6):
    for dx, dy in [(-4, -2), (3, -3), (4, 2), (0, 0), (2, -1), (5, -5), (-s, s), (s, -s)]:
        Bs = synth.crop(tex, 100 - dx, 100 - dy)
        fx, fy, q = R.compute_flow(A, Bs, s)
        shifts.append({"s": s, "dx": dx, "dy": dy, "fx": fx, "fy": fy, "q": q})
G["kat"]["lcg_shifts"] = shifts
Bh = ((B.astype(np.int32) + synth.crop(tex, 97, 101) + 1) // 2).astype(np.uint8)
G["kat"]["half_pixel"] = list(R.compute_flow(A, Bh, 4))
rng = np.random.default_rng(99)
low = (128 + rng.integers(0, 3, (64, 64))).astype(np.uint8)
G["kat"]["low_texture"] = list(R.compute_flow(low, np.roll(low, 1, axis=1), 4))

# bulk pairs: (seed, search, value threshold, n, max_shift) -> reference outputs
for seed, s, vt, n, ms in [(11, 4, 3000, 256, 4), (12, 4, 3000, 128, 6), (13, 6, 3000, 256, 6), (14, 8, 3000, 64, 8),
                           (15, 4, 900, 128, 4), (16, 6, 1200, 64, 7)]:
    prev, cur, _ = synth.sad_pairs(n, max_shift=ms, seed=seed, half_pixel_every=3)
    fx, fy, q = R.compute_flow_batch(prev, cur, s, 30, vt, threads=4)
    G["pairs"].append({"seed": seed, "search": s, "val_thr": vt, "n": n, "max_shift": ms, "half_pixel_every": 3,
                       "fx": [float(v) for v in fx], "fy": [float(v) for v in fy], "q": [int(v) for v in q]})

# calcFlow sequences (Appendix B.2): one reference instance at a time; the limiter's
# static time_last_pub is 0 in a fresh process, so these are generated in order and
# the first sequence sees t_last = 0.
def run_seq(rate, dt_step, nframes):
    o = R.px4(216.6677, 216.2457, rate, 64, 64, 4, 30, 3000)
    rows = []
    for t in range(nframes):
        img = synth.crop(tex, 100 - 2 * t, 100 + t)
        q, dt, fx, fy = o.calc_flow(img, 1000 + dt_step * t)
        rows.append({"q": q, "dt": dt, "fx": None if np.isnan(fx) else fx, "fy": None if np.isnan(fy) else fy})
    del o
    return rows

seq1 = run_seq(-1, 10000, 8)
# re-zero the static through a throw-away unlimited instance called at t=0 (dt wraps; ignored)
z = R.px4(1.0, 1.0, -1, 64, 64, 4, 30, 3000); z.calc_flow(A, 0); z.calc_flow(A, 0); del z
seq2 = run_seq(15, 13333, 20)
G["seq"] = [{"rate": -1, "dt_step": 10000, "rows": seq1}, {"rate": 15, "dt_step": 13333, "rows": seq2}]

with open(os.path.join(os.path.dirname(__file__), "sad_golden.json"), "w") as f:
    json.dump(G, f, indent=0)
print("wrote sad_golden.json:", {k: len(v) for k, v in G.items() if k != "source"})
print(G["kat"]["lcg_shifts"][:6], seq1[:3], [r for r in seq2 if r["q"] >= 0][:3])
