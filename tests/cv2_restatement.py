"""Literal Python restatement of the reference's OpenCV path over cv2 (test helper).

  OpticalFlowOpenCV::calcFlow  /root/reference/modules/OpticalFlow/src/flow_opencv.cpp:89-210
  trackFeatures/initMorePoints  .../external/klt_feature_tracker/src/trackFeatures.cpp:70-363
  OpticalFlow::limitRate        .../src/optical_flow.cpp:51-96

Every OpenCV call the reference makes is made here through cv2 (calcOpticalFlowPyrLK,
FAST via FastFeatureDetector).  State is per object.  The one defined divergence: the
response sort is stable (ties keep FAST's raster order), see SURVEY.md section 7.
Used to pin oracle/lk_oracle.c end to end and to generate tests/golden/lk_golden.json.
"""
import math

import numpy as np

try:
    import cv2
    cv2.setNumThreads(1)
except Exception:  # pragma: no cover
    cv2 = None

f32 = np.float32


class Cv2OpticalFlowOpenCV:
    def __init__(self, fx, fy, rate=15, w=64, h=64, n_feat=20, conf=1.645):
        self.fx, self.fy, self.rate, self.w, self.h, self.n, self.conf = f32(fx), f32(fy), rate, w, h, n_feat, f32(conf)
        self.update = None
        self.feat_prev = None
        # trackFeatures.cpp:53-56
        self.prev_img = None
        self.prev_corners = []          # list of [x, y] float32
        self.prev_status = [0] * 100
        # limiter
        self.sum_x = f32(0); self.sum_y = f32(0); self.sum_q = 0; self.n_valid = 0; self.t_last = 0
        self.det = cv2.FastFeatureDetector_create(10, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        self.fast_bins = 0

    # ---- trackFeatures.cpp:70-132
    def track_features(self, img, status):
        n = len(status)
        feats = [[f32(-100), f32(-100)] for _ in range(n)]
        for i in range(n):
            self.prev_status[i] = 1 if status[i] == 1 else 0
        if self.prev_img is not None and self.prev_corners:
            pts = np.array(self.prev_corners, np.float32).reshape(-1, 1, 2)
            cur, st, _ = cv2.calcOpticalFlowPyrLK(self.prev_img, img, pts, None, winSize=(9, 9), maxLevel=3)
            cur = cur.reshape(-1, 2); st = st.reshape(-1)
            self.prev_corners = [[f32(a), f32(b)] for a, b in cur]
            for i in range(min(len(self.prev_corners), n)):
                if not (self.prev_status[i] and st[i]):
                    self.prev_status[i] = 0
                if self.prev_status[i] == 1:
                    x, y = self.prev_corners[i]
                    if x < 0 or x > self.w or y < 0 or y > self.h:
                        status[i] = 0
                    else:
                        feats[i] = [x, y]
                        status[i] = 1
                elif status[i] == 1:
                    status[i] = 0
        self.prev_img = img.copy()
        self.init_more_points(img, status, feats)
        return feats

    def _close_existing(self, x, y, dist):
        for j in range(len(self.prev_corners)):
            if self.prev_status[j] == 0:
                continue
            ex, ey = int(self.prev_corners[j][0]), int(self.prev_corners[j][1])
            if abs(ex - x) < dist and abs(ey - y) < dist:
                return True
        return False

    # ---- trackFeatures.cpp:136-363 (stereo = 0)
    def init_more_points(self, img, update, feats):
        target = sum(1 for u in update if u == 2)
        self.fast_bins = 0
        if not target:
            return
        bw, bh = self.w // 4, self.h // 4
        tpb = (len(update) - 1) // 16 + 1
        per_bin = [[0] * 4 for _ in range(4)]
        for i in range(len(self.prev_corners)):
            if update[i] == 1:
                bx, by = int(self.prev_corners[i][0] / f32(bw)), int(self.prev_corners[i][1] / f32(bh))
                if bx >= 4 or by >= 4:
                    continue
                per_bin[bx][by] += 1
        dist = bw // tpb
        good, unused = [], []
        for x in range(4):
            for y in range(4):
                needed = max(0, tpb - per_bin[x][y])
                if not needed:
                    continue
                self.fast_bins += 1
                c0, c1, r0, r1 = x * bw, min((x + 1) * bw, self.w), y * bh, min((y + 1) * bh, self.h)
                kps = self.det.detect(np.ascontiguousarray(img[r0:r1, c0:c1]))
                kps = [(int(k.pt[0]) + c0, int(k.pt[1]) + r0, k.response) for k in kps]
                kps.sort(key=lambda k: -k[2])   # stable: ties keep raster order
                bin_good = []
                idx = 0
                broke = False
                while idx < len(kps):
                    nx, ny, _ = kps[idx]
                    far = not self._close_existing(nx, ny, dist)
                    if not far:
                        unused.append(kps[idx])
                    else:
                        for g in bin_good:
                            if abs(g[0] - nx) < dist and abs(g[1] - ny) < dist:
                                far = False
                                unused.append(kps[idx])
                                break
                        if far:
                            bin_good.append(kps[idx])
                            if len(bin_good) == needed:
                                broke = True
                                break
                    idx += 1
                good += bin_good
                size_minus_1 = len(kps) - 1 if len(kps) else (1 << 64) - 1   # size_t wrap
                if idx < size_minus_1:
                    unused += kps[idx:]
        if len(good) > target:
            n_remove = len(good) - target
            step = target // n_remove + 2
            good = [g for i, g in enumerate(good) if i % step]
        if len(good) < target:
            order = sorted(range(len(unused)), key=lambda i: (-unused[i][2], unused[i][1] * self.w + unused[i][0]))
            dist //= 2
            for i in order:
                nx, ny, _ = unused[i]
                if self._close_existing(nx, ny, dist):
                    continue
                if any(abs(g[0] - nx) < dist and abs(g[1] - ny) < dist for g in good):
                    continue
                good.append(unused[i])
                if len(good) == target:
                    break
        if not good:
            for i in range(len(update)):
                if update[i] == 2:
                    update[i] = 0
            return
        while len(self.prev_corners) < len(update):
            self.prev_corners.append([f32(0), f32(0)])
        m = 0
        for i in range(len(update)):
            if update[i] == 2:
                if m < len(good):
                    pt = [f32(good[m][0]), f32(good[m][1])]
                    self.prev_corners[i] = list(pt)
                    self.prev_status[i] = 1
                    feats[i] = list(pt)
                    m += 1
                else:
                    update[i] = 0

    # ---- optical_flow.cpp:51-96 (time_last_pub per object)
    def limit_rate(self, q, t_us, fx, fy):
        if self.rate <= 0:
            dt = (t_us - self.t_last) & 0xFFFFFFFF
            self.t_last = t_us
            return q, _as_int(dt), fx, fy
        if q > 0:
            self.sum_x = f32(self.sum_x + fx); self.sum_y = f32(self.sum_y + fy)
            self.sum_q += q; self.n_valid += 1
        if f32((t_us - self.t_last) & 0xFFFFFFFF) > f32(1.0e6) / f32(self.rate):
            avg = 0
            if self.n_valid > 0:
                avg = self.sum_q // self.n_valid   # int / int (both non-negative)
            fx, fy = self.sum_x, self.sum_y
            self.sum_x = f32(0); self.sum_y = f32(0); self.sum_q = 0; self.n_valid = 0
            dt = (t_us - self.t_last) & 0xFFFFFFFF
            self.t_last = t_us
            return avg, _as_int(dt), fx, fy
        return -1, None, fx, fy

    # ---- flow_opencv.cpp:89-210
    def calc_flow(self, img, t_us):
        if self.update is None:
            self.update = [2] * self.n
        feats = self.track_features(img, self.update)
        uv = self.update
        meancount = 0
        mx = my = f32(0)
        if self.feat_prev is not None:
            p = self.feat_prev
            for i in range(self.n):
                if uv[i] == 1:
                    mx = f32(mx + f32(feats[i][0] - p[i][0])); my = f32(my + f32(feats[i][1] - p[i][1]))
                    meancount += 1
            if meancount:
                mx = f32(mx / f32(meancount)); my = f32(my / f32(meancount))
                sx = sy = f32(0)
                for i in range(self.n):
                    if uv[i] == 1:
                        ex = f32(f32(feats[i][0] - p[i][0]) - mx); ey = f32(f32(feats[i][1] - p[i][1]) - my)
                        sx = f32(float(sx) + float(ex) * float(ex)); sy = f32(float(sy) + float(ey) * float(ey))
                sx = f32(np.sqrt(f32(sx / f32(meancount)))); sy = f32(np.sqrt(f32(sy / f32(meancount))))
                tx = ty = f32(0)
                meancount = 0
                for i in range(self.n):
                    if uv[i] == 1:
                        fx_ = f32(feats[i][0] - p[i][0]); fy_ = f32(feats[i][1] - p[i][1])
                        if abs(f32(fx_ - mx)) < f32(sx * self.conf) and abs(f32(fy_ - my)) < f32(sy * self.conf):
                            tx = f32(tx + fx_); ty = f32(ty + fy_); meancount += 1
                        else:
                            uv[i] = 0
                if meancount:
                    mx = f32(tx / f32(meancount)); my = f32(ty / f32(meancount))
        self.feat_prev = [list(f) for f in feats]
        for i in range(self.n):
            if uv[i] == 2:
                uv[i] = 1
            if uv[i] == 0:
                uv[i] = 2
        self.inliers = meancount
        self.px = (float(mx), float(my))
        q = int(_round_half_away(255.0 * meancount / self.n))
        q, dt, fx, fy = self.limit_rate(q, t_us, mx, my)
        return q, dt, float(f32(math.atan2(float(fx), float(self.fx)))), float(f32(math.atan2(float(fy), float(self.fy))))


def _as_int(u):
    return u - (1 << 32) if u >= (1 << 31) else u


def _round_half_away(x):
    return math.floor(x + 0.5) if x >= 0 else -math.floor(-x + 0.5)
