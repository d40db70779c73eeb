"""Golden vectors for the rows either side of the path, generated in the build container from
cv2.calcHist and from the reference's own MAVLink headers (oracle/_ref/libaof_ref_mavlink.so):
    python tests/golden/make_next_golden.py  ->  tests/golden/next_golden.json"""
import json
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import next_rows as nr  # noqa: E402

gold = {"ae": [], "flow_rad": []}
for k, (w, h) in enumerate([(640, 480), (128, 128), (200, 160), (1280, 720)]):
    rng = np.random.default_rng(100 + k)
    img = rng.integers(0, 256, (h, w), dtype=np.uint8)
    if k == 2:
        img[:] = np.clip(img.astype(int) + 120, 0, 255)  # many saturated (255) pixels: outside the histogram range
    mask = np.zeros((h, w), np.uint8)
    mask[h // 2 - 64:h // 2 + 64, w // 2 - 64:w // 2 + 64] = 255
    hist = cv2.calcHist([img], [0], mask, [10], [0, 255]).reshape(-1)
    msv = np.float32(0)
    for i in range(10):
        msv = np.float32(msv + np.float32((i + 1) * hist[i]) / np.float32(16384.0))
    gold["ae"].append({"w": w, "h": h, "seed": 100 + k, "bias": 120 if k == 2 else 0, "hist": [int(v) for v in hist], "msv": float(msv)})
rng = np.random.default_rng(7)
for k in range(12):
    c = {"time_usec": int(rng.integers(0, 2 ** 52)), "dt_us": int(rng.integers(0, 2 ** 31)), "flow_x": float(np.float32(rng.normal() * 0.02)),
         "flow_y": float(np.float32(rng.normal() * 0.02)), "gyro": [float(v) for v in rng.normal(size=3) * 0.1],
         "quality": int(rng.integers(0, 256)) if k % 4 else 0, "sysid": int(rng.integers(1, 255)), "compid": int(rng.integers(1, 255)), "seq": int(rng.integers(0, 256))}
    c["frame"] = nr.ref_flow_rad_frame(c["time_usec"], c["dt_us"], c["flow_x"], c["flow_y"], c["gyro"], c["quality"], c["sysid"], c["compid"], c["seq"]).hex()
    gold["flow_rad"].append(c)
json.dump(gold, open(os.path.join(ROOT, "tests", "golden", "next_golden.json"), "w"), indent=1)
print("wrote next_golden.json")
