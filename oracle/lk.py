"""oracle/lk.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes bindings for oracle/liboracle_lk.so (oracle/lk_oracle.c): the CPU restatement of
OpticalFlowOpenCV::calcFlow + trackFeatures + the OpenCV arithmetic they call.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "liboracle_lk.so")

u8p = C.POINTER(C.c_uint8)
f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int)
i16p = C.POINTER(C.c_int16)

_L = None


def lib():
    global _L
    if _L is not None:
        return _L
    if not os.path.exists(SO):
        subprocess.run(["make", "-s", "-C", HERE], check=True, stdout=subprocess.DEVNULL)
    L = C.CDLL(SO)
    L.olk_pyr_top_level.argtypes = [C.c_int] * 4
    L.olk_pyr_down.argtypes = [u8p, C.c_int, C.c_int, u8p]
    L.olk_pyr_down.restype = None
    L.olk_scharr.argtypes = [u8p, C.c_int, C.c_int, i16p]
    L.olk_scharr.restype = None
    L.olk_track.argtypes = [C.POINTER(u8p), C.POINTER(u8p), i32p, i32p, C.c_int, f32p, f32p, u8p, C.c_int, C.c_int,
                            C.c_int, C.c_double, C.c_float]
    L.olk_track.restype = None
    L.olk_fast.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, u8p]
    L.olk_new.restype = C.c_void_p
    L.olk_new.argtypes = [C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]
    L.olk_free.argtypes = [C.c_void_p]
    L.olk_free.restype = None
    L.olk_calc_flow.argtypes = [C.c_void_p, u8p, C.c_uint32, i32p, f32p, f32p]
    L.olk_get_state.argtypes = [C.c_void_p, i32p, f32p, f32p, u8p, i32p]
    L.olk_get_state.restype = None
    L.olk_set_camera_matrix.argtypes = [C.c_void_p] + [C.c_float] * 4
    L.olk_set_camera_matrix.restype = None
    L.olk_set_camera_distortion.argtypes = [C.c_void_p] + [C.c_float] * 5
    L.olk_set_camera_distortion.restype = None
    L.olk_set_params.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_int, C.c_float]
    L.olk_set_params.restype = None
    L.olk_undistort.argtypes = [f32p, f32p, f32p]
    L.olk_undistort.restype = None
    L.olk_min_eig.argtypes = [u8p, C.c_int, C.c_int, f32p]
    L.olk_min_eig.restype = None
    L.olk_gftt.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, f32p]
    L.olk_gftt.restype = C.c_int
    L.olk_set_sort_hook.argtypes = [C.c_void_p]
    L.olk_set_sort_hook.restype = None
    _L = L
    return L


def tie_sensitive_sorts(reset=True):
    """Sort calls since the last reset whose outcome depends on an unstable sort's tie order (> 16 elements with a tie)."""
    L = lib()
    L.olk_tie_sensitive_sorts.restype = C.c_long
    return L.olk_tie_sensitive_sorts(1 if reset else 0)


_STD = None


def use_std_sort(on):
    """Test hook: order response ties as libstdc++'s std::sort with the reference's comparator does
    (oracle/stdsort_shim.cpp) instead of the deterministic (response desc, raster asc) order.  Process-wide."""
    global _STD
    if on and _STD is None:
        so = os.path.join(HERE, "liboracle_stdsort.so")
        if not os.path.exists(so):
            subprocess.run(["make", "-s", "-C", HERE, so], check=True, stdout=subprocess.DEVNULL)
        _STD = C.CDLL(so)
    lib().olk_set_sort_hook(C.cast(_STD.olk_std_sort, C.c_void_p) if on else None)


def _p(a, t):
    return a.ctypes.data_as(t)


def pyr_top_level(w, h, win=9, max_level=3):
    return lib().olk_pyr_top_level(w, h, win, max_level)


def pyr_down(img):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.empty(((h + 1) // 2, (w + 1) // 2), np.uint8)
    lib().olk_pyr_down(_p(img, u8p), w, h, _p(out, u8p))
    return out


def build_pyramid(img, win=9, max_level=3):
    img = np.ascontiguousarray(img, np.uint8)
    top = pyr_top_level(img.shape[1], img.shape[0], win, max_level)
    pyr = [img]
    for _ in range(top):
        pyr.append(pyr_down(pyr[-1]))
    return pyr


def scharr(img):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.empty((h, w, 2), np.int16)
    lib().olk_scharr(_p(img, u8p), w, h, _p(out, i16p))
    return out


def track(pyr_prev, pyr_next, pts, win=9, max_iter=30, eps=0.01, min_eig=1e-4):
    """calcOpticalFlowPyrLK over prebuilt pyramids -> (next_pts [n,2] f32, status [n] u8)."""
    n_lv = len(pyr_prev)
    pp = (u8p * n_lv)(*[_p(a, u8p) for a in pyr_prev])
    pn = (u8p * n_lv)(*[_p(a, u8p) for a in pyr_next])
    ws = np.array([a.shape[1] for a in pyr_prev], np.int32)
    hs = np.array([a.shape[0] for a in pyr_prev], np.int32)
    pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
    n = len(pts)
    out = np.zeros((n, 2), np.float32)
    st = np.zeros(n, np.uint8)
    lib().olk_track(pp, pn, _p(ws, i32p), _p(hs, i32p), n_lv - 1, _p(pts, f32p), _p(out, f32p), _p(st, u8p), n, win,
                    max_iter, eps, min_eig)
    return out, st


def lk(prev, cur, pts, win=9, max_level=3):
    return track(build_pyramid(prev, win, max_level), build_pyramid(cur, win, max_level), pts, win)


def fast(roi, thr=10):
    """cv::FAST(roi, kps, thr) with NMS -> score map (0 = no keypoint, else response)."""
    roi = np.asarray(roi, np.uint8)
    h, w = roi.shape
    stride = roi.strides[0]
    assert roi.strides[1] == 1
    out = np.zeros((h, w), np.uint8)
    lib().olk_fast(C.cast(roi.ctypes.data, u8p), stride, w, h, thr, _p(out, u8p))
    return out


class Tracker:
    """One OpticalFlowOpenCV object (flow_opencv.hpp:81-83) with its own tracker state."""

    def __init__(self, fx, fy, rate=15, w=64, h=64, n_feat=20, conf=1.645):
        self.n = n_feat
        self.w, self.h = w, h
        self.h_ = lib().olk_new(fx, fy, rate, w, h, n_feat, conf)
        assert self.h_

    def __del__(self):
        if getattr(self, "h_", None):
            lib().olk_free(self.h_)
            self.h_ = None

    def set_camera_matrix(self, fx, fy, cx, cy):
        lib().olk_set_camera_matrix(self.h_, fx, fy, cx, cy)

    def set_camera_distortion(self, k1, k2, k3, p1=0.0, p2=0.0):
        lib().olk_set_camera_distortion(self.h_, k1, k2, k3, p1, p2)

    def set_params(self, fx, fy, rate, conf):
        lib().olk_set_params(self.h_, fx, fy, rate, conf)

    def calc_flow(self, img, t_us):
        img = np.ascontiguousarray(img, np.uint8)
        assert img.size == self.w * self.h
        dt = C.c_int(-123456)
        fx = C.c_float(np.nan)
        fy = C.c_float(np.nan)
        q = lib().olk_calc_flow(self.h_, _p(img, u8p), C.c_uint32(int(t_us)), C.byref(dt), C.byref(fx), C.byref(fy))
        return q, dt.value, fx.value, fy.value

    def state(self):
        uv = np.zeros(self.n, np.int32)
        fc = np.zeros((self.n, 2), np.float32)
        pc = np.zeros((self.n, 2), np.float32)
        ps = np.zeros(self.n, np.uint8)
        diag = np.zeros(8, np.int32)
        lib().olk_get_state(self.h_, _p(uv, i32p), _p(fc, f32p), _p(pc, f32p), _p(ps, u8p), _p(diag, i32p))
        return {"update_vec": uv, "feat_cur": fc, "prev_corners": pc, "prev_status": ps, "fast_bins": int(diag[0]),
                "inliers": int(diag[1]), "active": int(diag[2]), "n_prev_corners": int(diag[3]),
                "px_x": float(diag[4:5].view(np.float32)[0]), "px_y": float(diag[5:6].view(np.float32)[0])}


def undistort(cam, dc, pts):
    """cv::undistortPoints + re-projection of flow_opencv.cpp:108-119; cam=(fx,fy,cx,cy), dc=(k1,k2,p1,p2,k3)."""
    cam = np.ascontiguousarray(cam, np.float32)
    dc = np.ascontiguousarray(dc, np.float32)
    out = np.ascontiguousarray(pts, np.float32).reshape(-1, 2).copy()
    for i in range(len(out)):
        lib().olk_undistort(_p(cam, f32p), _p(dc, f32p), _p(out[i:i + 1], f32p))
    return out


def min_eig(img):
    """cv::cornerMinEigenVal(img, 3, 3)"""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    out = np.zeros((h, w), np.float32)
    lib().olk_min_eig(_p(img, u8p), w, h, _p(out, f32p))
    return out


def gftt(img, max_corners, quality, min_distance):
    """cv::goodFeaturesToTrack(img, max_corners, quality, min_distance) -> [n, 2] (x, y)"""
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    xy = np.zeros((max(1, max_corners), 2), np.float32)
    n = lib().olk_gftt(_p(img, u8p), w, h, max_corners, quality, min_distance, _p(xy, f32p))
    return xy[:n].copy()
