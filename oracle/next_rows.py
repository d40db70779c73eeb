"""oracle/next_rows.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatements of the steps either side of calcFlow, each citing the reference lines it follows
(paths relative to /root/reference/):

  crop_center      src/mainloop.cpp:294-298     cv::Rect(w/2 - cw/2, h/2 - ch/2, cw, ch) + copyTo (:317-319)
  ae_stats         src/mainloop.cpp:203-220     cv::calcHist(10 bins, range [0, 255), mask = central 128 x 128) + MSV
  flow_rad_frame   src/mainloop.cpp:359-371, src/mavlink_tcp.cpp:142-149, modules/mavlink_c/mavlink_helpers.h
                   (mavlink_finalize_message_buffer / mavlink_msg_to_send_buffer, MAVLink 2) and checksum.h (X.25)

Pinned by: cv2.calcHist (tests/test_next_rows.py, live) and the reference's own MAVLink headers compiled into
oracle/_ref/libaof_ref_mavlink.so (oracle/ref_mavlink_shim.cpp); golden vectors in tests/golden/next_golden.json.
"""
import ctypes as C
import os
import struct

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_MAVLINK_SO = os.path.join(HERE, "_ref", "libaof_ref_mavlink.so")


def crop_center(frame, w, h):
    H, W = frame.shape
    x0, y0 = W // 2 - w // 2, H // 2 - h // 2
    return np.ascontiguousarray(frame[y0:y0 + h, x0:x0 + w])


def ae_stats(frame):
    """-> (hist[10] uint32, msv float32)"""
    H, W = frame.shape
    x0, y0 = W // 2 - 64, H // 2 - 64
    roi = frame[y0:y0 + 128, x0:x0 + 128].astype(np.int64)
    roi = roi[roi < 255]                       # the range [0, 255) is half-open
    hist = np.bincount(roi * 10 // 255, minlength=10).astype(np.uint32)
    msv = np.float32(0.0)
    for i in range(10):                        # float accumulation in the reference's order
        msv = np.float32(msv + np.float32(np.float32(i + 1) * np.float32(hist[i])) / np.float32(16384.0))
    return hist, msv


def _crc_acc(b, crc):
    tmp = (b ^ (crc & 0xFF)) & 0xFF
    tmp = (tmp ^ (tmp << 4)) & 0xFF
    return ((crc >> 8) ^ (tmp << 8) ^ (tmp << 3) ^ (tmp >> 4)) & 0xFFFF


def flow_rad_frame(time_usec, dt_us, flow_x, flow_y, gyro, quality, system_id, component_id, seq):
    payload = struct.pack("<QIfffffIfhBB", time_usec & 0xFFFFFFFFFFFFFFFF, dt_us & 0xFFFFFFFF, np.float32(flow_x), np.float32(flow_y),
                          np.float32(-gyro[1]), np.float32(gyro[0]), np.float32(gyro[2]), 0, -1.0, 0, 0, quality & 0xFF)
    assert len(payload) == 44
    n = len(payload)
    while n > 1 and payload[n - 1] == 0:
        n -= 1
    payload = payload[:n]
    hdr = bytes([0xFD, n, 0, 0, seq & 0xFF, system_id, component_id, 106, 0, 0])
    crc = 0xFFFF
    for b in hdr[1:] + payload:
        crc = _crc_acc(b, crc)
    crc = _crc_acc(138, crc)
    return hdr + payload + bytes([crc & 0xFF, crc >> 8])


def ref_mavlink_available():
    return os.path.exists(REF_MAVLINK_SO)


def ref_flow_rad_frame(time_usec, dt_us, flow_x, flow_y, gyro, quality, system_id, component_id, seq):
    """the reference's own MAVLink headers (oracle/_ref/libaof_ref_mavlink.so)"""
    L = C.CDLL(REF_MAVLINK_SO)
    L.ref_flow_rad_frame.argtypes = [C.c_uint64, C.c_int, C.c_float, C.c_float, C.c_double, C.c_double, C.c_double, C.c_int,
                                     C.c_uint8, C.c_uint8, C.c_uint8, C.c_void_p]
    buf = (C.c_uint8 * 300)()
    n = L.ref_flow_rad_frame(time_usec, dt_us, flow_x, flow_y, gyro[0], gyro[1], gyro[2], quality, system_id, component_id, seq, buf)
    return bytes(buf[:n])
