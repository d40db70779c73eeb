"""Generates tests/golden/lk_golden.json from cv2 4.13 (the oracle of record for the OpenCV
arithmetic) and from tests/cv2_restatement.py (the reference's
trackFeatures + flow_opencv logic driven through cv2), and -- "seq_ref" -- from the reference's own LK objects
(oracle/_ref/libaof_ref_lk.so) with cv2 behind them.  Run in the authoring container:
    python tests/golden/make_lk_golden.py
The fixture pins oracle/lk_oracle.c on machines without cv2."""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import cv2  # noqa: E402

import __graft_entry__ as entry  # noqa: E402
from cv2_restatement import Cv2OpticalFlowOpenCV  # noqa: E402

cv2.setNumThreads(1)
pkg = entry.load_package()
FX, FY = 216.6677, 216.2457


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()[:16]


out = {"cv2": cv2.__version__, "pyr": [], "lk": [], "fast": [], "seq": []}
# pyramids: hash of every level
for (w, h, seed) in [(640, 480, 0), (128, 128, 1), (64, 64, 2), (75, 61, 3)]:
    img = pkg.synth.lk_sequence(1, w, h, 1, 1, seed=seed)[0]
    lv, a = [], img
    for _ in range(3):
        a = cv2.pyrDown(a)
        lv.append(sha(a))
    out["pyr"].append({"w": w, "h": h, "seed": seed, "levels": lv})
# LK on fixed points (incl. border / outside cases)
rng = np.random.default_rng(11)
for (w, h, seed, dx, dy) in [(640, 480, 0, -3, 2), (128, 128, 1, 2, -1), (64, 64, 2, 1, 0)]:
    fr = pkg.synth.lk_sequence(2, w, h, dx, dy, seed=seed)
    pts = np.concatenate([rng.uniform([0, 0], [w, h], (40, 2)),
                          [[0, 0], [w - 1, h - 1], [w - 2, h / 2], [w / 2, 0.5], [1.5, h - 3], [-3, h / 3], [w + 9, 5]]]).astype(np.float32)
    c, st, _ = cv2.calcOpticalFlowPyrLK(fr[0], fr[1], pts.reshape(-1, 1, 2), None, winSize=(9, 9), maxLevel=3)
    out["lk"].append({"w": w, "h": h, "seed": seed, "dx": dx, "dy": dy, "pts": pts.tolist(),
                      "next": c.reshape(-1, 2).astype(np.float64).tolist(), "status": st.reshape(-1).tolist()})
# FAST per ROI: keypoint count + hash of the score map
det = cv2.FastFeatureDetector_create(10, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
img = pkg.synth.lk_sequence(1, 640, 480, 1, 1, seed=0)[0]
for (y0, x0, hh, ww) in [(0, 0, 120, 160), (120, 320, 120, 160), (0, 0, 32, 32), (50, 60, 37, 41), (360, 480, 120, 160)]:
    roi = np.ascontiguousarray(img[y0:y0 + hh, x0:x0 + ww])
    m = np.zeros((hh, ww), np.uint8)
    kps = det.detect(roi)
    for k in kps:
        m[int(k.pt[1]), int(k.pt[0])] = int(k.response)
    out["fast"].append({"roi": [y0, x0, hh, ww], "n": len(kps), "sha": sha(m)})
# calcFlow sequences
for (w, h, dx, dy, seed, rate, nf, T) in [(640, 480, -3, 2, 0, -1, 20, 10), (640, 480, -3, 2, 1, 15, 20, 12),
                                           (128, 128, 2, -1, 2, -1, 20, 10), (128, 128, 1, 1, 3, 15, 20, 14),
                                           (64, 64, 1, 0, 4, -1, 20, 8), (320, 240, -5, 4, 5, -1, 33, 8)]:
    fr = pkg.synth.lk_sequence(T, w, h, dx, dy, seed=seed)
    B = Cv2OpticalFlowOpenCV(FX, FY, rate, w, h, nf)
    rows = []
    for k in range(T):
        q, dt, fx, fy = B.calc_flow(fr[k], 1000 + 13333 * k)
        rows.append({"q": q, "dt": dt, "fx": fx, "fy": fy, "inliers": B.inliers, "px": list(B.px),
                     "update": list(B.update), "fast_bins": B.fast_bins})
    out["seq"].append({"w": w, "h": h, "dx": dx, "dy": dy, "seed": seed, "rate": rate, "nf": nf, "rows": rows})
# calcFlow sequences from the REFERENCE'S OWN OBJECTS (oracle/_ref/libaof_ref_lk.so: trackFeatures.cpp, flow_opencv.cpp,
# optical_flow.cpp compiled unmodified) with cv2 executing behind their three OpenCV calls.  These carry the
# reference's real std::sort tie order; tests/test_ref_lk.py replays them on the oracle with libstdc++'s sort installed.
from oracle import ref_lk  # noqa: E402

out["seq_ref"] = []
for (w, h, dx, dy, seed, rate, nf, T) in [(640, 480, -3, 2, 0, -1, 20, 10), (640, 480, -3, 2, 1, 15, 20, 12),
                                           (128, 128, 2, -1, 2, -1, 20, 10), (128, 128, 1, 1, 3, 15, 20, 14),
                                           (64, 64, 1, 0, 4, -1, 20, 8), (320, 240, -5, 4, 5, -1, 33, 8)]:
    fr = pkg.synth.lk_sequence(T, w, h, dx, dy, seed=seed)
    R = ref_lk.RefTracker(FX, FY, rate, w, h, nf, backend="cv2")
    rows = []
    for k in range(T):
        q, dt, fx, fy = R.calc_flow(fr[k], 1000 + 13333 * k)
        st = R.state()
        rows.append({"q": q, "dt": dt, "fx": fx, "fy": fy, "update": st["update_vec"].tolist(),
                     "feat": st["feat_cur"].astype(np.float64).tolist()})
    R.close()
    out["seq_ref"].append({"w": w, "h": h, "dx": dx, "dy": dy, "seed": seed, "rate": rate, "nf": nf, "rows": rows})
json.dump(out, open(os.path.join(HERE, "lk_golden.json"), "w"))
print("wrote lk_golden.json", os.path.getsize(os.path.join(HERE, "lk_golden.json")))
