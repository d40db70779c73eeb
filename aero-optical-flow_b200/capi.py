"""ctypes view of the C ABI in include/aof.h (libaof_b200.so).

This is the binding a Python caller (tests, bench.py) uses; there is no other
implementation behind it: if the library is missing or no CUDA device is
present the calls raise, they never fall back to a CPU path.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libaof_b200.so")

AOF_OK = 0
AOF_ERR_BASE = -100
AOF_PX4_MAX_BLOCKS = 25


class AofError(RuntimeError):
    pass


class Px4Params(C.Structure):
    _fields_ = [("focal_length_x", C.c_float), ("focal_length_y", C.c_float), ("output_rate", C.c_int),
                ("image_width", C.c_int), ("image_height", C.c_int), ("search_size", C.c_int),
                ("flow_feature_threshold", C.c_int), ("flow_value_threshold", C.c_int)]


flow_out_dtype = np.dtype([("quality", np.int32), ("dt_us", np.int32), ("flow_x", np.float32),
                           ("flow_y", np.float32)])
px4_raw_dtype = np.dtype([("px_x", np.float32), ("px_y", np.float32), ("quality", np.int32), ("count", np.int32)])
px4_block_dtype = np.dtype([("x", np.int16), ("y", np.int16), ("dx", np.int16), ("dy", np.int16),
                            ("texture", np.uint32), ("best_sad", np.uint32), ("acc", np.uint32, (8,)),
                            ("subdir", np.int32), ("accepted", np.int32)])

class LkParams(C.Structure):
    _fields_ = [("focal_length_x", C.c_float), ("focal_length_y", C.c_float), ("output_rate", C.c_int),
                ("image_width", C.c_int), ("image_height", C.c_int), ("num_features", C.c_int),
                ("confidence_multiplier", C.c_float)]


lk_raw_dtype = np.dtype([("px_x", np.float32), ("px_y", np.float32), ("inliers", np.int32), ("active", np.int32),
                         ("fast_bins", np.int32), ("quality", np.int32)])

class DenseParams(C.Structure):
    _fields_ = [("image_width", C.c_int), ("image_height", C.c_int), ("search_size", C.c_int),
                ("flow_feature_threshold", C.c_int), ("flow_value_threshold", C.c_int)]


ae_stats_dtype = np.dtype([("hist", np.uint32, (10,)), ("msv", np.float32), ("pad", np.uint32)])

dense_block_dtype = np.dtype([("dx", np.int16), ("dy", np.int16), ("subdir", np.int16), ("flags", np.int16),
                              ("best_sad", np.uint32), ("texture", np.uint32)])

_lib = None


def lib():
    """Load libaof_b200.so (built in-tree by build.sh / __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise AofError(f"{LIB_PATH} not built: run __graft_entry__.build() (no CPU fallback exists)")
    L = C.CDLL(LIB_PATH)
    vp, u8p, u32p, ip, fp = C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_float)
    L.aof_version.restype = C.c_char_p
    L.aof_last_error.restype = C.c_char_p
    L.aof_device_count.restype = C.c_int
    L.aof_launch_count.restype = C.c_uint64
    L.aof_timing_enable.argtypes = [C.c_int]
    L.aof_timing_enable.restype = None
    L.aof_timing_reset.restype = None
    L.aof_timing_read.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
    L.aof_px4_default_params.argtypes = [C.POINTER(Px4Params)]
    L.aof_px4_create.argtypes = [C.POINTER(Px4Params), C.c_int, C.c_int, C.POINTER(vp)]
    L.aof_px4_destroy.argtypes = [vp]
    L.aof_px4_destroy.restype = None
    L.aof_px4_num_streams.argtypes = [vp]
    L.aof_px4_reset.argtypes = [vp]
    L.aof_px4_calc_flow.argtypes = [vp, u8p, C.c_uint32, ip, fp, fp]
    L.aof_px4_calc_flow_batch.argtypes = [vp, u8p, u32p, vp]
    L.aof_px4_calc_flow_device.argtypes = [vp, u8p, u32p, vp, vp, vp, vp]
    L.aof_px4_compute_flow_pairs.argtypes = [vp, u8p, u8p, C.c_int, vp, vp]
    L.aof_px4_compute_flow_pairs_device.argtypes = [vp, u8p, u8p, C.c_int, vp, vp, vp]
    L.aof_lk_default_params.argtypes = [C.POINTER(LkParams)]
    L.aof_lk_create.argtypes = [C.POINTER(LkParams), C.c_int, C.c_int, C.POINTER(vp)]
    L.aof_lk_destroy.argtypes = [vp]
    L.aof_lk_destroy.restype = None
    L.aof_lk_num_streams.argtypes = [vp]
    L.aof_lk_reset.argtypes = [vp]
    L.aof_lk_calc_flow.argtypes = [vp, u8p, C.c_uint32, ip, fp, fp]
    L.aof_lk_calc_flow_batch.argtypes = [vp, u8p, u32p, vp, vp]
    L.aof_lk_calc_flow_device.argtypes = [vp, u8p, u32p, vp, vp, vp]
    L.aof_lk_get_state.argtypes = [vp, vp, vp, vp, vp]
    L.aof_px4_set_params.argtypes = [vp, C.c_float, C.c_float, C.c_int]
    L.aof_lk_set_params.argtypes = [vp, C.c_float, C.c_float, C.c_int, C.c_float]
    L.aof_lk_set_camera_matrix.argtypes = [vp] + [C.c_float] * 4
    L.aof_lk_set_camera_distortion.argtypes = [vp] + [C.c_float] * 5
    L.aof_lk_pyramid.argtypes = [vp, u8p, C.c_int, vp, vp, vp]
    L.aof_lk_track_points.argtypes = [vp, u8p, u8p, C.c_int, vp, C.c_int, vp, vp]
    L.aof_lk_fast_scores.argtypes = [vp, u8p, C.c_int, vp]
    L.aof_sad_dense_create.argtypes = [C.POINTER(DenseParams), C.c_int, C.POINTER(vp)]
    L.aof_sad_dense_destroy.argtypes = [vp]
    L.aof_sad_dense_destroy.restype = None
    L.aof_sad_dense_grid.argtypes = [vp, ip, ip, ip, ip]
    L.aof_sad_dense_compute.argtypes = [vp, u8p, u8p, C.c_int, vp]
    L.aof_sad_dense_compute_device.argtypes = [vp, u8p, u8p, C.c_int, vp, vp]
    L.aof_gftt_create.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    L.aof_gftt_destroy.argtypes = [vp]
    L.aof_gftt_destroy.restype = None
    L.aof_gftt_min_eig.argtypes = [vp, u8p, C.c_int, vp]
    L.aof_gftt_detect.argtypes = [vp, u8p, C.c_int, C.c_int, C.c_double, C.c_double, vp, vp]
    L.aof_lk_lazy_stats.argtypes = [vp, C.POINTER(C.c_uint64)]
    L.aof_lk_num_features.argtypes = [vp]
    L.aof_px4_create_multi.argtypes = [C.POINTER(Px4Params), C.c_int, ip, C.c_int, C.POINTER(vp)]
    L.aof_px4_destroy_multi.argtypes = [vp]
    L.aof_px4_destroy_multi.restype = None
    L.aof_px4_multi_calc_flow_batch.argtypes = [vp, u8p, u32p, vp]
    L.aof_px4_multi_results.argtypes = [vp]
    L.aof_px4_multi_results.restype = vp
    L.aof_px4_multi_range.argtypes = [vp, C.c_int, ip, ip, ip]
    L.aof_px4_multi_set_params.argtypes = [vp, C.c_float, C.c_float, C.c_int]
    L.aof_lk_create_multi.argtypes = [C.POINTER(LkParams), C.c_int, ip, C.c_int, C.POINTER(vp)]
    L.aof_lk_destroy_multi.argtypes = [vp]
    L.aof_lk_destroy_multi.restype = None
    L.aof_lk_multi_calc_flow_batch.argtypes = [vp, u8p, u32p, vp, vp]
    L.aof_lk_multi_results.argtypes = [vp]
    L.aof_lk_multi_results.restype = vp
    L.aof_lk_multi_raw.argtypes = [vp]
    L.aof_lk_multi_raw.restype = vp
    L.aof_lk_multi_range.argtypes = [vp, C.c_int, ip, ip, ip]
    L.aof_lk_multi_set_params.argtypes = [vp, C.c_float, C.c_float, C.c_int, C.c_float]
    L.aof_lk_multi_get_state.argtypes = [vp, vp, vp, vp, vp]
    L.aof_host_register.argtypes = [vp, C.c_size_t]
    L.aof_host_unregister.argtypes = [vp]
    L.aof_crop_to_device.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
    L.aof_ae_stats_device.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp]
    L.aof_ae_stats_host.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]
    L.aof_pack_optical_flow_rad.argtypes = [vp, C.c_int, C.c_uint64, vp, vp, C.c_uint8, C.c_uint8, vp, vp, C.c_int, vp]
    _lib = L
    return L


def kernel_times():
    """{kernel name: (total ms, launches)} recorded since aof_timing_reset (needs aof_timing_enable(1))."""
    L = lib()
    out = {}
    for k in range(L.aof_timing_kernels()):
        name = C.c_char_p(); ms = C.c_double(); cnt = C.c_uint64()
        L.aof_timing_read(k, C.byref(name), C.byref(ms), C.byref(cnt))
        if cnt.value:
            out[name.value.decode()] = (ms.value, cnt.value)
    return out


def check(rc, what=""):
    if rc <= AOF_ERR_BASE:
        raise AofError(f"{what} failed ({rc}): {lib().aof_last_error().decode()}")
    return rc


def _ptr(a):
    return C.c_void_p(a.ctypes.data)


class Px4Flow:
    """n_streams OpticalFlowPX4 instances on one GPU (mirrors flow_px4.hpp:62-68)."""

    def __init__(self, focal_x, focal_y, output_rate=15, width=64, height=64, search=6, feat_thr=30, val_thr=3000,
                 n_streams=1, device=0):
        L = lib()
        self.params = Px4Params(focal_x, focal_y, output_rate, width, height, search, feat_thr, val_thr)
        self.n, self.w, self.h = n_streams, width, height
        self.handle = C.c_void_p()
        check(L.aof_px4_create(C.byref(self.params), n_streams, device, C.byref(self.handle)), "aof_px4_create")

    def close(self):
        if getattr(self, "handle", None) and lib is not None:   # (module globals are gone at interpreter exit)
            lib().aof_px4_destroy(self.handle)
            self.handle = None

    __del__ = close

    def reset(self):
        check(lib().aof_px4_reset(self.handle), "aof_px4_reset")

    def calc_flow(self, img, t_us):
        """One stream.  -> (quality, dt_us, flow_x, flow_y); outputs are None where untouched."""
        img = np.ascontiguousarray(img, np.uint8)
        assert img.size == self.w * self.h
        dt = C.c_int(-123456); fx = C.c_float(np.nan); fy = C.c_float(np.nan)
        q = check(lib().aof_px4_calc_flow(self.handle, _ptr(img), t_us, C.byref(dt), C.byref(fx), C.byref(fy)),
                  "aof_px4_calc_flow")
        return q, dt.value, fx.value, fy.value

    def calc_flow_batch(self, imgs, t_us):
        imgs = np.ascontiguousarray(imgs, np.uint8)
        t_us = np.ascontiguousarray(t_us, np.uint32)
        assert imgs.size == self.n * self.w * self.h and t_us.size == self.n
        out = np.zeros(self.n, flow_out_dtype)
        check(lib().aof_px4_calc_flow_batch(self.handle, _ptr(imgs), _ptr(t_us), _ptr(out)), "aof_px4_calc_flow_batch")
        return out

    def set_params(self, focal_x, focal_y, output_rate):
        """setFocalLengthX/Y, setOutputRate on the live handle: previous frames and limiter state are kept."""
        check(lib().aof_px4_set_params(self.handle, focal_x, focal_y, output_rate), "aof_px4_set_params")

    def calc_flow_device(self, d_imgs, d_t_us, d_out, d_raw=0, d_blocks=0, stream=0):
        """Raw device pointers (ints); asynchronous on `stream`."""
        check(lib().aof_px4_calc_flow_device(self.handle, d_imgs, d_t_us, d_out, d_raw or None, d_blocks or None,
                                             stream or None), "aof_px4_calc_flow_device")

    def compute_flow_pairs(self, prev, cur, want_blocks=False):
        prev = np.ascontiguousarray(prev, np.uint8); cur = np.ascontiguousarray(cur, np.uint8)
        n = prev.size // (self.w * self.h)
        raw = np.zeros(n, px4_raw_dtype)
        blocks = np.zeros((n, AOF_PX4_MAX_BLOCKS), px4_block_dtype) if want_blocks else None
        check(lib().aof_px4_compute_flow_pairs(self.handle, _ptr(prev), _ptr(cur), n, _ptr(raw),
                                               _ptr(blocks) if want_blocks else None), "aof_px4_compute_flow_pairs")
        return (raw, blocks) if want_blocks else raw

    def compute_flow_pairs_device(self, d_prev, d_cur, n, d_raw, d_blocks=0, stream=0):
        check(lib().aof_px4_compute_flow_pairs_device(self.handle, d_prev, d_cur, n, d_raw, d_blocks or None,
                                                      stream or None), "aof_px4_compute_flow_pairs_device")


class LkFlow:
    """n_streams OpticalFlowOpenCV instances on one GPU (mirrors flow_opencv.hpp:73-88)."""

    def __init__(self, focal_x, focal_y, output_rate=15, width=64, height=64, num_features=20, conf_multi=1.645,
                 n_streams=1, device=0):
        L = lib()
        self.params = LkParams(focal_x, focal_y, output_rate, width, height, num_features, conf_multi)
        self.n, self.w, self.h, self.nf = n_streams, width, height, num_features
        self.handle = C.c_void_p()
        check(L.aof_lk_create(C.byref(self.params), n_streams, device, C.byref(self.handle)), "aof_lk_create")

    def close(self):
        if getattr(self, "handle", None) and lib is not None:   # (module globals are gone at interpreter exit)
            lib().aof_lk_destroy(self.handle)
            self.handle = None

    __del__ = close

    def reset(self):
        check(lib().aof_lk_reset(self.handle), "aof_lk_reset")

    def calc_flow(self, img, t_us):
        img = np.ascontiguousarray(img, np.uint8)
        assert img.size == self.w * self.h
        dt = C.c_int(-123456); fx = C.c_float(np.nan); fy = C.c_float(np.nan)
        q = check(lib().aof_lk_calc_flow(self.handle, _ptr(img), int(t_us), C.byref(dt), C.byref(fx), C.byref(fy)),
                  "aof_lk_calc_flow")
        return q, dt.value, fx.value, fy.value

    def calc_flow_batch(self, imgs, t_us, want_raw=False):
        imgs = np.ascontiguousarray(imgs, np.uint8)
        t_us = np.ascontiguousarray(t_us, np.uint32)
        assert imgs.size == self.n * self.w * self.h and t_us.size == self.n
        out = np.zeros(self.n, flow_out_dtype)
        raw = np.zeros(self.n, lk_raw_dtype) if want_raw else None
        check(lib().aof_lk_calc_flow_batch(self.handle, _ptr(imgs), _ptr(t_us), _ptr(out),
                                           _ptr(raw) if want_raw else None), "aof_lk_calc_flow_batch")
        return (out, raw) if want_raw else out

    def set_params(self, focal_x, focal_y, output_rate, conf_multi):
        """setFocalLengthX/Y, setOutputRate, setConfMultiplier on the live handle: all state is kept."""
        check(lib().aof_lk_set_params(self.handle, focal_x, focal_y, output_rate, conf_multi), "aof_lk_set_params")

    def lazy_stats(self):
        """(bins picked lazily, pool rebuilds, bins that overflowed, stream-frames sent to the complete path)"""
        out = (C.c_uint64 * 4)()
        check(lib().aof_lk_lazy_stats(self.handle, out), "aof_lk_lazy_stats")
        return tuple(int(v) for v in out)

    def set_camera_matrix(self, fx, fy, cx, cy):
        check(lib().aof_lk_set_camera_matrix(self.handle, fx, fy, cx, cy), "aof_lk_set_camera_matrix")

    def set_camera_distortion(self, k1, k2, k3, p1=0.0, p2=0.0):
        check(lib().aof_lk_set_camera_distortion(self.handle, k1, k2, k3, p1, p2), "aof_lk_set_camera_distortion")

    def calc_flow_device(self, d_imgs, d_t_us, d_out, d_raw=0, stream=0):
        check(lib().aof_lk_calc_flow_device(self.handle, d_imgs, d_t_us, d_out, d_raw or None, stream or None),
              "aof_lk_calc_flow_device")

    def state(self):
        uv = np.zeros((self.n, self.nf), np.int32)
        fc = np.zeros((self.n, self.nf, 2), np.float32)
        pc = np.zeros((self.n, self.nf, 2), np.float32)
        ps = np.zeros((self.n, self.nf), np.uint8)
        check(lib().aof_lk_get_state(self.handle, _ptr(uv), _ptr(fc), _ptr(pc), _ptr(ps)), "aof_lk_get_state")
        return {"update_vec": uv, "feat_cur": fc, "prev_corners": pc, "prev_status": ps}

    def pyramid(self, imgs):
        imgs = np.ascontiguousarray(imgs, np.uint8).reshape(-1, self.h, self.w)
        n = len(imgs)
        lw = np.zeros(4, np.int32); lh = np.zeros(4, np.int32)
        nl = check(lib().aof_lk_pyramid(self.handle, None, 0, _ptr(lw), _ptr(lh), None), "aof_lk_pyramid")
        per = int(sum(int(lw[l]) * int(lh[l]) for l in range(nl)))
        buf = np.zeros((n, per), np.uint8)
        check(lib().aof_lk_pyramid(self.handle, _ptr(imgs), n, _ptr(lw), _ptr(lh), _ptr(buf)), "aof_lk_pyramid")
        levels, o = [], 0
        for l in range(nl):
            sz = int(lw[l]) * int(lh[l])
            levels.append(buf[:, o:o + sz].reshape(n, int(lh[l]), int(lw[l])))
            o += sz
        return levels

    def track_points(self, prev, cur, pts):
        prev = np.ascontiguousarray(prev, np.uint8).reshape(-1, self.h, self.w)
        cur = np.ascontiguousarray(cur, np.uint8).reshape(-1, self.h, self.w)
        n = len(prev)
        pts = np.ascontiguousarray(pts, np.float32).reshape(n, -1, 2)
        out = np.zeros_like(pts)
        st = np.zeros(pts.shape[:2], np.uint8)
        check(lib().aof_lk_track_points(self.handle, _ptr(prev), _ptr(cur), n, _ptr(pts), pts.shape[1], _ptr(out),
                                        _ptr(st)), "aof_lk_track_points")
        return out, st

    def fast_scores(self, imgs):
        imgs = np.ascontiguousarray(imgs, np.uint8).reshape(-1, self.h, self.w)
        sc = np.zeros_like(imgs)
        check(lib().aof_lk_fast_scores(self.handle, _ptr(imgs), len(imgs), _ptr(sc)), "aof_lk_fast_scores")
        return sc


class _Multi:
    """Shared part of the multi-GPU wrappers: whole streams per device, one host worker thread per device inside the
    library, results gathered into ONE pinned host array (no collective)."""
    _kind = None

    def _devs(self, devices):
        devices = list(range(lib().aof_device_count())) if devices is None else list(devices)
        return devices, (C.c_int * len(devices))(*devices)

    def close(self):
        if getattr(self, "handle", None) and lib is not None:   # (module globals are gone at interpreter exit)
            getattr(lib(), "aof_%s_destroy_multi" % self._kind)(self.handle)
            self.handle = None

    __del__ = close

    def ranges(self):
        """[(device, first stream, stream count)] -- the shard of each device"""
        out = []
        for i in range(self.n_devices):
            d, f, c = C.c_int(), C.c_int(), C.c_int()
            check(getattr(lib(), "aof_%s_multi_range" % self._kind)(self.handle, i, C.byref(d), C.byref(f), C.byref(c)), "multi_range")
            out.append((d.value, f.value, c.value))
        return out

    def _view(self, addr, dtype):
        buf = (C.c_char * (self.n * dtype.itemsize)).from_address(addr)
        return np.frombuffer(buf, dtype=dtype, count=self.n)


class Px4FlowMulti(_Multi):
    """n_streams OpticalFlowPX4 instances sharded over `devices` (default: all) -- aof_px4_create_multi."""
    _kind = "px4"

    def __init__(self, focal_x, focal_y, output_rate=15, width=64, height=64, search=6, feat_thr=30, val_thr=3000,
                 n_streams=1, devices=None):
        L = lib()
        self.params = Px4Params(focal_x, focal_y, output_rate, width, height, search, feat_thr, val_thr)
        self.n, self.w, self.h = n_streams, width, height
        devs, arr = self._devs(devices)
        self.n_devices = len(devs)
        self.handle = C.c_void_p()
        check(L.aof_px4_create_multi(C.byref(self.params), n_streams, arr, len(devs), C.byref(self.handle)), "aof_px4_create_multi")

    def calc_flow_batch(self, imgs, t_us, copy=True):
        """imgs: n_streams frames in host memory (pinned for full speed).  copy=False returns a view of the library's
        pinned result array (valid until the next call)."""
        imgs = np.ascontiguousarray(imgs, np.uint8)
        t_us = np.ascontiguousarray(t_us, np.uint32)
        assert imgs.size == self.n * self.w * self.h and t_us.size == self.n
        check(lib().aof_px4_multi_calc_flow_batch(self.handle, _ptr(imgs), _ptr(t_us), None), "aof_px4_multi_calc_flow_batch")
        v = self._view(lib().aof_px4_multi_results(self.handle), flow_out_dtype)
        return v.copy() if copy else v

    def set_params(self, focal_x, focal_y, output_rate):
        check(lib().aof_px4_multi_set_params(self.handle, focal_x, focal_y, output_rate), "aof_px4_multi_set_params")


class LkFlowMulti(_Multi):
    """n_streams OpticalFlowOpenCV instances sharded over `devices` (default: all) -- aof_lk_create_multi."""
    _kind = "lk"

    def __init__(self, focal_x, focal_y, output_rate=15, width=64, height=64, num_features=20, conf_multi=1.645,
                 n_streams=1, devices=None):
        L = lib()
        self.params = LkParams(focal_x, focal_y, output_rate, width, height, num_features, conf_multi)
        self.n, self.w, self.h, self.nf = n_streams, width, height, num_features
        devs, arr = self._devs(devices)
        self.n_devices = len(devs)
        self.handle = C.c_void_p()
        check(L.aof_lk_create_multi(C.byref(self.params), n_streams, arr, len(devs), C.byref(self.handle)), "aof_lk_create_multi")

    def calc_flow_batch(self, imgs, t_us, want_raw=False, copy=True):
        imgs = np.ascontiguousarray(imgs, np.uint8)
        t_us = np.ascontiguousarray(t_us, np.uint32)
        assert imgs.size == self.n * self.w * self.h and t_us.size == self.n
        L = lib()
        raw_addr = L.aof_lk_multi_raw(self.handle)
        check(L.aof_lk_multi_calc_flow_batch(self.handle, _ptr(imgs), _ptr(t_us), None, raw_addr if want_raw else None),
              "aof_lk_multi_calc_flow_batch")
        out = self._view(L.aof_lk_multi_results(self.handle), flow_out_dtype)
        out = out.copy() if copy else out
        if not want_raw:
            return out
        raw = self._view(raw_addr, lk_raw_dtype)
        return out, (raw.copy() if copy else raw)

    def set_params(self, focal_x, focal_y, output_rate, conf_multi):
        check(lib().aof_lk_multi_set_params(self.handle, focal_x, focal_y, output_rate, conf_multi), "aof_lk_multi_set_params")

    def state(self):
        uv = np.zeros((self.n, self.nf), np.int32)
        fc = np.zeros((self.n, self.nf, 2), np.float32)
        pc = np.zeros((self.n, self.nf, 2), np.float32)
        ps = np.zeros((self.n, self.nf), np.uint8)
        check(lib().aof_lk_multi_get_state(self.handle, _ptr(uv), _ptr(fc), _ptr(pc), _ptr(ps)), "aof_lk_multi_get_state")
        return {"update_vec": uv, "feat_cur": fc, "prev_corners": pc, "prev_status": ps}


class SadDense:
    """dense per-block SAD flow field (aof_sad_dense_*)."""

    def __init__(self, width, height, search=8, feat_thr=30, val_thr=3000, device=0):
        self.params = DenseParams(width, height, search, feat_thr, val_thr)
        self.w, self.h = width, height
        self.handle = C.c_void_p()
        check(lib().aof_sad_dense_create(C.byref(self.params), device, C.byref(self.handle)), "aof_sad_dense_create")
        g = [C.c_int() for _ in range(4)]
        lib().aof_sad_dense_grid(self.handle, *[C.byref(x) for x in g])
        self.nbx, self.nby, self.x0, self.y0 = [x.value for x in g]

    def close(self):
        if getattr(self, "handle", None) and lib is not None:   # (module globals are gone at interpreter exit)
            lib().aof_sad_dense_destroy(self.handle)
            self.handle = None

    __del__ = close

    def compute(self, prev, cur):
        prev = np.ascontiguousarray(prev, np.uint8).reshape(-1, self.h, self.w)
        cur = np.ascontiguousarray(cur, np.uint8).reshape(-1, self.h, self.w)
        n = len(prev)
        out = np.zeros((n, self.nby, self.nbx), dense_block_dtype)
        check(lib().aof_sad_dense_compute(self.handle, _ptr(prev), _ptr(cur), n, _ptr(out)), "aof_sad_dense_compute")
        return out

    def compute_device(self, d_prev, d_cur, n, d_field, stream=0):
        check(lib().aof_sad_dense_compute_device(self.handle, d_prev, d_cur, n, d_field, stream or None),
              "aof_sad_dense_compute_device")


class Gftt:
    """cv::cornerMinEigenVal / cv::goodFeaturesToTrack on the GPU (aof_gftt_*)."""

    def __init__(self, width, height, device=0):
        self.w, self.h = width, height
        self.handle = C.c_void_p()
        check(lib().aof_gftt_create(width, height, device, C.byref(self.handle)), "aof_gftt_create")

    def close(self):
        if getattr(self, "handle", None) and lib is not None:   # (module globals are gone at interpreter exit)
            lib().aof_gftt_destroy(self.handle)
            self.handle = None

    __del__ = close

    def min_eig(self, imgs):
        imgs = np.ascontiguousarray(imgs, np.uint8).reshape(-1, self.h, self.w)
        out = np.zeros(imgs.shape, np.float32)
        check(lib().aof_gftt_min_eig(self.handle, _ptr(imgs), len(imgs), _ptr(out)), "aof_gftt_min_eig")
        return out

    def detect(self, imgs, max_corners, quality, min_distance):
        imgs = np.ascontiguousarray(imgs, np.uint8).reshape(-1, self.h, self.w)
        n = len(imgs)
        xy = np.zeros((n, max_corners, 2), np.float32)
        cnt = np.zeros(n, np.int32)
        check(lib().aof_gftt_detect(self.handle, _ptr(imgs), n, max_corners, quality, min_distance, _ptr(xy), _ptr(cnt)),
              "aof_gftt_detect")
        return [xy[i, :cnt[i]] for i in range(n)]


# ---- the steps either side of calcFlow (include/aof.h "callers / consumers") ----------------

def ae_stats(frames, device=0):
    """auto-exposure statistics (src/mainloop.cpp:197-220) of n host frames [n, h, w] -> ae_stats_dtype[n]"""
    frames = np.ascontiguousarray(frames, np.uint8)
    n, h, w = frames.shape
    out = np.zeros(n, ae_stats_dtype)
    check(lib().aof_ae_stats_host(_ptr(frames), n, w, h, device, _ptr(out)), "aof_ae_stats_host")
    return out


def pack_optical_flow_rad(res, time_offset_us, img_time_us, gyro_xyz, system_id, component_id, seq):
    """MAVLink 2 OPTICAL_FLOW_RAD frames of the published results (quality >= 0).
    -> (bytes, frame offsets[n] (-1 = not published), next seq)"""
    res = np.ascontiguousarray(res, flow_out_dtype)
    n = len(res)
    t = np.ascontiguousarray(img_time_us, np.uint32)
    g = np.ascontiguousarray(gyro_xyz, np.float64).reshape(n, 3)
    buf = np.zeros(56 * max(n, 1), np.uint8)
    off = np.zeros(max(n, 1), np.int32)
    sq = np.array([seq], np.uint8)
    used = check(lib().aof_pack_optical_flow_rad(_ptr(res), n, time_offset_us, _ptr(t), _ptr(g), system_id, component_id,
                                                 _ptr(sq), _ptr(buf), len(buf), _ptr(off)), "aof_pack_optical_flow_rad")
    return buf[:used].tobytes(), off[:n], int(sq[0])


def crop_to_device(cam_frames, w, h, d_batch_ptr, stream=0):
    """centre crops of host camera frames [n, cam_h, cam_w] -> contiguous device batch (strided H2D DMA, no host copy)"""
    cam_frames = np.ascontiguousarray(cam_frames, np.uint8)
    n, ch, cw = cam_frames.shape
    ptrs = (C.c_void_p * n)(*[cam_frames[i].ctypes.data for i in range(n)])
    check(lib().aof_crop_to_device(ptrs, n, cw, ch, w, h, C.c_void_p(d_batch_ptr), C.c_void_p(stream)), "aof_crop_to_device")
    return cam_frames  # keep alive until the stream has been synchronised
