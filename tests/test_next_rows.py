"""The rows either side of calcFlow that SURVEY.md 8(f) ranks "next": centre crop + staging, auto-exposure statistics,
OPTICAL_FLOW_RAD wire format.  CPU tests pin the oracle (oracle/next_rows.py) to the golden vectors (cv2.calcHist and the
reference's own MAVLink headers) and check the host-side packer of the library; GPU tests check the device paths."""
import json
import os

import numpy as np
import pytest

from oracle import next_rows as nr

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "next_golden.json")))


def ae_image(c):
    rng = np.random.default_rng(c["seed"])
    img = rng.integers(0, 256, (c["h"], c["w"]), dtype=np.uint8)
    if c["bias"]:
        img[:] = np.clip(img.astype(int) + c["bias"], 0, 255)
    return img


def test_oracle_ae_stats_golden():
    for c in GOLD["ae"]:
        hist, msv = nr.ae_stats(ae_image(c))
        assert hist.tolist() == c["hist"] and float(msv) == c["msv"]


def test_oracle_ae_stats_vs_cv2_live():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    for k in range(6):
        img = rng.integers(0, 256, (300, 260 + 8 * k), dtype=np.uint8)
        h, w = img.shape
        mask = np.zeros((h, w), np.uint8)
        mask[h // 2 - 64:h // 2 + 64, w // 2 - 64:w // 2 + 64] = 255
        want = cv2.calcHist([img], [0], mask, [10], [0, 255]).reshape(-1)
        hist, _ = nr.ae_stats(img)
        assert np.array_equal(hist.astype(np.float32), want)


def test_oracle_flow_rad_golden_and_reference_headers():
    for c in GOLD["flow_rad"]:
        f = nr.flow_rad_frame(c["time_usec"], c["dt_us"], c["flow_x"], c["flow_y"], c["gyro"], c["quality"], c["sysid"], c["compid"], c["seq"])
        assert f.hex() == c["frame"]
        if nr.ref_mavlink_available():
            assert f == nr.ref_flow_rad_frame(c["time_usec"], c["dt_us"], c["flow_x"], c["flow_y"], c["gyro"], c["quality"],
                                              c["sysid"], c["compid"], c["seq"])


def test_library_flow_rad_packer(pkg):
    """host-side entry point: no GPU needed.  One batch with published and unpublished (-1) streams."""
    cs = GOLD["flow_rad"]
    n = len(cs) + 2
    res = np.zeros(n, pkg.capi.flow_out_dtype)
    t = np.zeros(n, np.uint32)
    gyro = np.zeros((n, 3))
    off0 = 1234567890123
    want, k = [], 0
    seq0 = 250                                     # wraps through 255 -> 0
    for i in range(n):
        if i in (3, 7):
            res[i]["quality"] = -1                 # not published: no frame, no sequence number
            want.append(None)
            continue
        c = cs[k]; k += 1
        res[i] = (c["quality"], c["dt_us"], c["flow_x"], c["flow_y"])
        t[i] = (c["time_usec"] - off0) & 0xFFFFFFFF
        gyro[i] = c["gyro"]
        want.append(nr.flow_rad_frame(off0 + int(t[i]), c["dt_us"], c["flow_x"], c["flow_y"], c["gyro"], c["quality"], 1, 42,
                                      (seq0 + len([w for w in want if w is not None])) & 0xFF))
    buf, off, seq = pkg.capi.pack_optical_flow_rad(res, off0, t, gyro, 1, 42, seq0)
    assert seq == (seq0 + len(cs)) & 0xFF
    pos = 0
    for i in range(n):
        if want[i] is None:
            assert off[i] == -1
            continue
        assert off[i] == pos and buf[pos:pos + len(want[i])] == want[i], i
        pos += len(want[i])
    assert pos == len(buf)
    # too small a buffer is an error, not a truncation
    L = pkg.lib()
    small = np.zeros(20, np.uint8)
    sq = np.array([0], np.uint8)
    rc = L.aof_pack_optical_flow_rad(res.ctypes.data, n, 0, t.ctypes.data, gyro.ctypes.data, 1, 1, sq.ctypes.data, small.ctypes.data, 20, None)
    assert rc == -100


@pytest.mark.gpu
def test_ae_stats_gpu_matches_oracle(pkg):
    for c in GOLD["ae"]:
        img = ae_image(c)
        out = pkg.capi.ae_stats(img[None])
        assert out["hist"][0].tolist() == c["hist"] and float(out["msv"][0]) == c["msv"]
    rng = np.random.default_rng(11)
    imgs = rng.integers(0, 256, (5, 480, 640), dtype=np.uint8)
    imgs[1] = 255
    imgs[2] = 0
    out = pkg.capi.ae_stats(imgs)
    for i in range(5):
        hist, msv = nr.ae_stats(imgs[i])
        assert np.array_equal(out["hist"][i], hist) and out["msv"][i] == msv
    with pytest.raises(pkg.AofError):
        pkg.capi.ae_stats(np.zeros((1, 64, 64), np.uint8))   # the window is 128 x 128


@pytest.mark.gpu
def test_crop_to_device_then_calcflow(pkg):
    """camera frames (registered host memory) -> strided DMA crop -> calcFlow on the device batch == calcFlow on host crops"""
    import torch
    rng = np.random.default_rng(5)
    n, cw, ch, w, h = 6, 320, 240, 64, 64
    from oracle import sad as osad
    prev_c, cur_c, _ = pkg.synth.sad_pairs(n, max_shift=4, seed=21, half_pixel_every=3)
    cams = rng.integers(0, 256, (2, n, ch, cw), dtype=np.uint8)
    x0, y0 = cw // 2 - w // 2, ch // 2 - h // 2
    cams[0, :, y0:y0 + h, x0:x0 + w] = prev_c
    cams[1, :, y0:y0 + h, x0:x0 + w] = cur_c
    for k in range(2):
        for i in range(n):
            assert np.array_equal(nr.crop_center(cams[k, i], w, h), (prev_c, cur_c)[k][i])
    L = pkg.lib()
    assert L.aof_host_register(cams.ctypes.data, cams.nbytes) == 0
    try:
        d = torch.empty((2, n, h, w), dtype=torch.uint8, device="cuda")
        d_t = torch.zeros((2, n), dtype=torch.int32, device="cuda")
        d_t[1] += 5000
        d_out = torch.zeros((n, 4), dtype=torch.int32, device="cuda")
        flow = pkg.Px4Flow(200.0, 200.0, -1, w, h, 4, 30, 3000, n_streams=n)
        st = torch.cuda.current_stream().cuda_stream
        for k in range(2):
            pkg.capi.crop_to_device(cams[k], w, h, d[k].data_ptr(), st)
            flow.calc_flow_device(d[k].data_ptr(), d_t[k].data_ptr(), d_out.data_ptr(), stream=st)
        torch.cuda.synchronize()
        assert np.array_equal(d[1].cpu().numpy(), cur_c)
        got = d_out.cpu().numpy().view(pkg.capi.flow_out_dtype).reshape(-1)
        ref = pkg.Px4Flow(200.0, 200.0, -1, w, h, 4, 30, 3000, n_streams=n)
        ref.calc_flow_batch(prev_c, np.zeros(n, np.uint32))
        want = ref.calc_flow_batch(cur_c, np.full(n, 5000, np.uint32))
        for f in ("quality", "dt_us", "flow_x", "flow_y"):
            assert np.array_equal(got[f], want[f]), f
        ex, ey, eq = osad.Port().compute_flow_batch(prev_c, cur_c, 4)
        assert np.array_equal(got["quality"], eq)
        flow.close(); ref.close()
    finally:
        assert L.aof_host_unregister(cams.ctypes.data) == 0
    # bad geometry
    assert L.aof_crop_to_device(None, 1, 64, 64, 128, 128, None, None) == -100
