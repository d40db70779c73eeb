"""oracle/ref_lk.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes bindings for oracle/_ref/libaof_ref_lk.so: the reference's OWN LK-path objects
(trackFeatures.cpp, flow_opencv.cpp, optical_flow.cpp compiled unmodified against oracle/cv_compat/,
see oracle/Makefile and oracle/ref_lk_shim.cpp).

The reference keeps tracker state in file statics (trackFeatures.cpp:53-56) and the limiter's
time_last_pub in a function static (optical_flow.cpp:55), so one loaded library = one tracker:
`RefTracker` dlopens a private copy of the .so (fresh statics) and removes the copy at once.

Backends of the three OpenCV calls (cv_compat.hpp):
  "oracle" (default) -- oracle/lk_oracle.c's cv2-pinned primitives, linked into the library;
  "cv2"              -- the real OpenCV through the cv2 module (callbacks), when cv2 is importable.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
import ctypes as C
import os
import shutil
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_ref", "libaof_ref_lk.so")

u8p = C.POINTER(C.c_uint8)
f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int)

FAST_FN = C.CFUNCTYPE(C.c_int, u8p, C.c_int, C.c_int, C.c_int, C.c_int, f32p, C.c_int)
PYRLK_FN = C.CFUNCTYPE(None, u8p, u8p, C.c_int, C.c_int, f32p, C.c_int, f32p, u8p, C.c_int, C.c_int)
UNDIST_FN = C.CFUNCTYPE(None, f32p, C.c_int, f32p, f32p, f32p)


def available():
    return os.path.exists(SO)


def _load_private_copy():
    d = tempfile.mkdtemp(prefix="aof_ref_lk_")
    try:
        dst = os.path.join(d, "libaof_ref_lk.so")
        shutil.copyfile(SO, dst)
        L = C.CDLL(dst, mode=os.RTLD_LOCAL | os.RTLD_NOW)
    finally:
        shutil.rmtree(d, ignore_errors=True)     # the mapping stays valid after unlink
    L.ref_lk_new.restype = C.c_void_p
    L.ref_lk_new.argtypes = [C.c_float, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float]
    L.ref_lk_free.argtypes = [C.c_void_p]
    L.ref_lk_free.restype = None
    L.ref_lk_calc_flow.argtypes = [C.c_void_p, u8p, C.c_uint32, i32p, f32p, f32p]
    L.ref_lk_get_state.argtypes = [C.c_void_p, i32p, f32p]
    L.ref_lk_get_state.restype = None
    L.ref_lk_set_camera_matrix.argtypes = [C.c_void_p] + [C.c_float] * 4
    L.ref_lk_set_camera_matrix.restype = None
    L.ref_lk_set_camera_distortion.argtypes = [C.c_void_p] + [C.c_float] * 5
    L.ref_lk_set_camera_distortion.restype = None
    L.ref_lk_set_params.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_int, C.c_float]
    L.ref_lk_set_params.restype = None
    L.ref_lk_set_backend.argtypes = [FAST_FN, PYRLK_FN, UNDIST_FN]
    L.ref_lk_set_backend.restype = None
    L.ref_lk_call_counts.argtypes = [C.POINTER(C.c_long)]
    L.ref_lk_call_counts.restype = None
    return L


def _cv2_backend():
    """The three OpenCV calls of the path made through cv2 with the reference's own arguments."""
    import cv2
    cv2.setNumThreads(1)
    det = cv2.FastFeatureDetector_create(10, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)

    def fast(roi, stride, w, h, thr, xyr, cap):
        a = np.ctypeslib.as_array(roi, shape=((h - 1) * stride + w,))
        img = np.lib.stride_tricks.as_strided(a, shape=(h, w), strides=(stride, 1))
        if thr != det.getThreshold():
            det.setThreshold(thr)
        kps = det.detect(np.ascontiguousarray(img))
        n = min(len(kps), cap)
        if n:
            out = np.ctypeslib.as_array(xyr, shape=(cap * 3,))[: 3 * n].reshape(n, 3)
            out[:, :2] = cv2.KeyPoint_convert(kps[:n])
            out[:, 2] = np.fromiter((k.response for k in kps[:n]), np.float32, n)
        return n

    def pyrlk(prev, nxt, w, h, pts, n, out, status, win, max_level):
        p = np.ctypeslib.as_array(prev, shape=(h, w))
        q = np.ctypeslib.as_array(nxt, shape=(h, w))
        src = np.ctypeslib.as_array(pts, shape=(n, 2)).copy().reshape(-1, 1, 2)
        nx, st, _ = cv2.calcOpticalFlowPyrLK(p, q, src, None, winSize=(win, win), maxLevel=max_level)
        np.ctypeslib.as_array(out, shape=(n, 2))[:] = nx.reshape(-1, 2)
        np.ctypeslib.as_array(status, shape=(n,))[:] = st.reshape(-1)

    def undist(pts, n, K9, D5, out):
        src = np.ctypeslib.as_array(pts, shape=(n, 2)).copy().reshape(-1, 1, 2)
        K = np.ctypeslib.as_array(K9, shape=(3, 3)).copy()
        D = np.ctypeslib.as_array(D5, shape=(1, 5)).copy()
        np.ctypeslib.as_array(out, shape=(n, 2))[:] = cv2.undistortPoints(src, K, D).reshape(-1, 2)

    return FAST_FN(fast), PYRLK_FN(pyrlk), UNDIST_FN(undist)


class RefTracker:
    """One OpticalFlowOpenCV object of the reference (flow_opencv.hpp:81-83), in a private copy of the library."""

    def __init__(self, fx, fy, rate=15, w=64, h=64, n_feat=20, conf=1.645, backend="oracle"):
        self.L = _load_private_copy()
        self.n, self.w, self.h = n_feat, w, h
        self._cb = None
        if backend == "cv2":
            self._cb = _cv2_backend()          # keep the callbacks alive
            self.L.ref_lk_set_backend(*self._cb)
        elif backend != "oracle":
            raise ValueError(backend)
        self.h_ = self.L.ref_lk_new(fx, fy, rate, w, h, n_feat, conf)

    def close(self):
        if getattr(self, "h_", None):
            self.L.ref_lk_free(self.h_)
            self.h_ = None

    __del__ = close

    def set_camera_matrix(self, fx, fy, cx, cy):
        self.L.ref_lk_set_camera_matrix(self.h_, fx, fy, cx, cy)

    def set_camera_distortion(self, k1, k2, k3, p1=0.0, p2=0.0):
        self.L.ref_lk_set_camera_distortion(self.h_, k1, k2, k3, p1, p2)

    def set_params(self, fx, fy, rate, conf):
        self.L.ref_lk_set_params(self.h_, fx, fy, rate, conf)

    def calc_flow(self, img, t_us):
        img = np.ascontiguousarray(img, np.uint8)
        assert img.size == self.w * self.h
        dt = C.c_int(-123456)
        fx = C.c_float(np.nan)
        fy = C.c_float(np.nan)
        q = self.L.ref_lk_calc_flow(self.h_, img.ctypes.data_as(u8p), C.c_uint32(int(t_us)), C.byref(dt), C.byref(fx),
                                    C.byref(fy))
        return q, dt.value, fx.value, fy.value

    def state(self):
        uv = np.zeros(self.n, np.int32)
        fc = np.zeros((self.n, 2), np.float32)
        self.L.ref_lk_get_state(self.h_, uv.ctypes.data_as(i32p), fc.ctypes.data_as(f32p))
        return {"update_vec": uv, "feat_cur": fc}

    def call_counts(self):
        c = (C.c_long * 3)()
        self.L.ref_lk_call_counts(c)
        return {"fast": c[0], "pyrlk": c[1], "undistort": c[2]}
