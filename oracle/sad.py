"""oracle/sad.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes bindings for the two SAD checkers:
  * ``Port``  -> oracle/liboracle_sad.so  (our C restatement, oracle/sad_oracle.c)
  * ``Ref``   -> oracle/_ref/libaof_ref.so (the unmodified reference objects + ref_shim.cpp)
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "liboracle_sad.so")
REF_SO = os.path.join(HERE, "_ref", "libaof_ref.so")

u8p = C.POINTER(C.c_uint8)
f32p = C.POINTER(C.c_float)
i32p = C.POINTER(C.c_int)
u32p = C.POINTER(C.c_uint32)


def build():
    """(Re)build the checkers; the reference objects only when /root/reference exists."""
    subprocess.run(["make", "-s", "-C", HERE], check=True, stdout=subprocess.DEVNULL)


def _p(a, t):
    return a.ctypes.data_as(t)


class BlockInfo(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32), ("texture", C.c_uint32), ("searched", C.c_int32),
                ("best_sad", C.c_uint32), ("dx", C.c_int32), ("dy", C.c_int32), ("accepted", C.c_int32),
                ("acc", C.c_uint32 * 8), ("subdir", C.c_int32)]


class Port:
    """Our restatement (always available after build())."""

    def __init__(self):
        if not os.path.exists(PORT_SO):
            build()
        L = self.L = C.CDLL(PORT_SO)
        L.orc_block_texture.restype = C.c_uint32
        L.orc_block_texture.argtypes = [u8p, C.c_int, C.c_int, C.c_int]
        L.orc_sad8x8.restype = C.c_uint32
        L.orc_sad8x8.argtypes = [u8p, u8p] + [C.c_int] * 5
        L.orc_subpixel.restype = None
        L.orc_subpixel.argtypes = [u8p, u8p] + [C.c_int] * 5 + [u32p]
        L.orc_compute_flow.restype = C.c_int
        L.orc_compute_flow.argtypes = [u8p, u8p, C.c_int, C.c_int, C.c_uint32, C.c_uint32, f32p, f32p,
                                       C.POINTER(BlockInfo)]
        L.orc_compute_flow_batch.restype = None
        L.orc_compute_flow_batch.argtypes = [u8p, u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint32,
                                             f32p, f32p, i32p]
        L.orc_dense_grid.restype = None
        L.orc_dense_grid.argtypes = [C.c_int] * 3 + [i32p] * 4
        L.orc_dense_field.restype = None
        L.orc_dense_field.argtypes = [u8p, u8p, C.c_int, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p]
        L.orc_px4_new.restype = C.c_void_p
        L.orc_px4_new.argtypes = [C.c_float, C.c_float] + [C.c_int] * 6
        L.orc_px4_free.argtypes = [C.c_void_p]
        L.orc_px4_calc_flow.restype = C.c_int
        L.orc_px4_calc_flow.argtypes = [C.c_void_p, u8p, C.c_uint32, i32p, f32p, f32p, f32p, f32p, i32p]

    def texture(self, img, x, y):
        img = np.ascontiguousarray(img, np.uint8)
        return self.L.orc_block_texture(_p(img, u8p), x, y, img.shape[1])

    def sad8x8(self, a, b, ax, ay, bx, by):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return self.L.orc_sad8x8(_p(a, u8p), _p(b, u8p), ax, ay, bx, by, a.shape[1])

    def subpixel(self, a, b, ax, ay, bx, by):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        acc = np.zeros(8, np.uint32)
        self.L.orc_subpixel(_p(a, u8p), _p(b, u8p), ax, ay, bx, by, a.shape[1], _p(acc, u32p))
        return acc

    def compute_flow(self, prev, cur, search=4, feat_thr=30, val_thr=3000, want_blocks=False):
        prev = np.ascontiguousarray(prev, np.uint8); cur = np.ascontiguousarray(cur, np.uint8)
        fx, fy = C.c_float(0), C.c_float(0)
        info = (BlockInfo * 25)() if want_blocks else None
        q = self.L.orc_compute_flow(_p(prev, u8p), _p(cur, u8p), prev.shape[1], search, feat_thr, val_thr,
                                    C.byref(fx), C.byref(fy), info)
        if want_blocks:
            return fx.value, fy.value, q, info
        return fx.value, fy.value, q

    def compute_flow_batch(self, prev, cur, search=4, feat_thr=30, val_thr=3000):
        prev = np.ascontiguousarray(prev, np.uint8); cur = np.ascontiguousarray(cur, np.uint8)
        n, h, w = prev.shape
        fx = np.zeros(n, np.float32); fy = np.zeros(n, np.float32); q = np.zeros(n, np.int32)
        self.L.orc_compute_flow_batch(_p(prev, u8p), _p(cur, u8p), n, w, h, search, feat_thr, val_thr,
                                      _p(fx, f32p), _p(fy, f32p), _p(q, i32p))
        return fx, fy, q

    def px4(self, fx, fy, rate=15, w=64, h=64, search=6, feat_thr=30, val_thr=3000):
        return _PortPX4(self.L, fx, fy, rate, w, h, search, feat_thr, val_thr)


class _PortPX4:
    def __init__(self, L, *a):
        self.L = L
        self.h = L.orc_px4_new(*a)

    def calc_flow(self, img, t_us):
        """-> (quality, dt_us|None, flow_x|None, flow_y|None, px_x, px_y, raw_q); None = untouched."""
        img = np.ascontiguousarray(img, np.uint8)
        dt = C.c_int(-123456); fx = C.c_float(np.nan); fy = C.c_float(np.nan)
        px = C.c_float(np.nan); py = C.c_float(np.nan); rq = C.c_int(-2)
        q = self.L.orc_px4_calc_flow(self.h, _p(img, u8p), t_us, C.byref(dt), C.byref(fx), C.byref(fy),
                                     C.byref(px), C.byref(py), C.byref(rq))
        return q, dt.value, fx.value, fy.value, px.value, py.value, rq.value

    def set_params(self, fx, fy, rate):
        self.L.orc_px4_set_params.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_int]
        self.L.orc_px4_set_params.restype = None
        self.L.orc_px4_set_params(self.h, fx, fy, rate)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_px4_free(self.h)
            self.h = None


def ref_available():
    return os.path.exists(REF_SO)


class Ref:
    """The unmodified reference objects (prebuilt oracle/_ref/libaof_ref.so)."""

    def __init__(self):
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(REF_SO + " missing: run `make -C oracle` where /root/reference exists")
        L = self.L = C.CDLL(REF_SO)
        L.ref_compute_diff.restype = C.c_uint32
        L.ref_compute_diff.argtypes = [u8p, C.c_int, C.c_int, C.c_int]
        L.ref_compute_sad_8x8.restype = C.c_uint32
        L.ref_compute_sad_8x8.argtypes = [u8p, u8p] + [C.c_int] * 5
        L.ref_compute_subpixel.restype = None
        L.ref_compute_subpixel.argtypes = [u8p, u8p] + [C.c_int] * 5 + [u32p]
        L.ref_compute_flow.restype = C.c_int
        L.ref_compute_flow.argtypes = [u8p, u8p] + [C.c_int] * 4 + [f32p, f32p]
        L.ref_compute_flow_batch.restype = None
        L.ref_compute_flow_batch.argtypes = [u8p, u8p] + [C.c_int] * 6 + [f32p, f32p, i32p, C.c_int]
        L.ref_px4_new.restype = C.c_void_p
        L.ref_px4_new.argtypes = [C.c_float, C.c_float] + [C.c_int] * 6
        L.ref_px4_free.argtypes = [C.c_void_p]
        L.ref_px4_calc_flow.restype = C.c_int
        L.ref_px4_calc_flow.argtypes = [C.c_void_p, u8p, C.c_uint32, i32p, f32p, f32p]

    def texture(self, img, x, y):
        img = np.ascontiguousarray(img, np.uint8)
        return self.L.ref_compute_diff(_p(img, u8p), x, y, img.shape[1])

    def sad8x8(self, a, b, ax, ay, bx, by):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return self.L.ref_compute_sad_8x8(_p(a, u8p), _p(b, u8p), ax, ay, bx, by, a.shape[1])

    def subpixel(self, a, b, ax, ay, bx, by):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        acc = np.zeros(8, np.uint32)
        self.L.ref_compute_subpixel(_p(a, u8p), _p(b, u8p), ax, ay, bx, by, a.shape[1], _p(acc, u32p))
        return acc

    def compute_flow(self, prev, cur, search=4, feat_thr=30, val_thr=3000):
        prev = np.ascontiguousarray(prev, np.uint8); cur = np.ascontiguousarray(cur, np.uint8)
        fx, fy = C.c_float(0), C.c_float(0)
        q = self.L.ref_compute_flow(_p(prev, u8p), _p(cur, u8p), prev.shape[1], search, feat_thr, val_thr,
                                    C.byref(fx), C.byref(fy))
        return fx.value, fy.value, q

    def compute_flow_batch(self, prev, cur, search=4, feat_thr=30, val_thr=3000, threads=1):
        prev = np.ascontiguousarray(prev, np.uint8); cur = np.ascontiguousarray(cur, np.uint8)
        n, h, w = prev.shape
        fx = np.zeros(n, np.float32); fy = np.zeros(n, np.float32); q = np.zeros(n, np.int32)
        self.L.ref_compute_flow_batch(_p(prev, u8p), _p(cur, u8p), n, w, h, search, feat_thr, val_thr,
                                      _p(fx, f32p), _p(fy, f32p), _p(q, i32p), threads)
        return fx, fy, q

    def px4(self, fx, fy, rate=15, w=64, h=64, search=6, feat_thr=30, val_thr=3000):
        return _RefPX4(self.L, fx, fy, rate, w, h, search, feat_thr, val_thr)


class _RefPX4:
    """One live instance at a time (limitRate's static, optical_flow.cpp:55)."""

    def __init__(self, L, *a):
        self.L = L
        self.h = L.ref_px4_new(*a)

    def calc_flow(self, img, t_us):
        img = np.ascontiguousarray(img, np.uint8)
        dt = C.c_int(-123456); fx = C.c_float(np.nan); fy = C.c_float(np.nan)
        q = self.L.ref_px4_calc_flow(self.h, _p(img, u8p), t_us, C.byref(dt), C.byref(fx), C.byref(fy))
        return q, dt.value, fx.value, fy.value

    def set_params(self, fx, fy, rate):
        self.L.ref_px4_set_params.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_int]
        self.L.ref_px4_set_params.restype = None
        self.L.ref_px4_set_params(self.h, fx, fy, rate)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_px4_free(self.h)
            self.h = None


dense_block_dtype = np.dtype([("dx", np.int16), ("dy", np.int16), ("subdir", np.int16), ("flags", np.int16),
                              ("best_sad", np.uint32), ("texture", np.uint32)])


def dense_field(port, prev, cur, search=8, feat_thr=30, val_thr=3000):
    """oracle dense field of one frame pair -> array [nby, nbx] of dense_block_dtype"""
    prev = np.ascontiguousarray(prev, np.uint8); cur = np.ascontiguousarray(cur, np.uint8)
    h, w = prev.shape
    g = [C.c_int() for _ in range(4)]
    port.L.orc_dense_grid(w, h, search, *[C.byref(x) for x in g])
    nbx, nby = g[2].value, g[3].value
    out = np.zeros((nby, nbx), dense_block_dtype)
    port.L.orc_dense_field(_p(prev, u8p), _p(cur, u8p), w, h, search, feat_thr, val_thr, C.c_void_p(out.ctypes.data))
    return out
