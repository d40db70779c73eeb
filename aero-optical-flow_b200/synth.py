"""Deterministic synthetic translated-texture frames.

Pure numpy; used by tests, smoke() and bench.py to build inputs.  Nothing here is
on the measured path.
"""
import numpy as np


def lcg_texture(size=256, seed=12345):
    """Appendix B.1 texture: LCG bytes, 3x3 wrap-around box sum // 9."""
    s = seed & 0xFFFFFFFF
    raw = np.empty(size * size, np.uint8)
    for i in range(size * size):
        s = (s * 1664525 + 1013904223) & 0xFFFFFFFF
        raw[i] = (s >> 8) & 0xFF
    raw = raw.reshape(size, size).astype(np.int32)
    acc = np.zeros_like(raw)
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            acc += np.roll(np.roll(raw, -dy, axis=0), -dx, axis=1)
    return (acc // 9).astype(np.uint8)


def crop(tex, ox, oy, w=64, h=64):
    """frame(ox, oy): w x h crop with its top-left at column ox, row oy."""
    return np.ascontiguousarray(tex[oy:oy + h, ox:ox + w])


def box_blur3(a):
    """3x3 box (integer //9) over the last two axes, edge-replicated."""
    p = np.pad(a.astype(np.int32), [(0, 0)] * (a.ndim - 2) + [(1, 1), (1, 1)], mode="edge")
    acc = np.zeros(a.shape, np.int32)
    H, W = a.shape[-2:]
    for dy in range(3):
        for dx in range(3):
            acc += p[..., dy:dy + H, dx:dx + W]
    return (acc // 9).astype(np.uint8)


def sad_pairs(n, w=64, h=64, max_shift=4, seed=0, half_pixel_every=0):
    """n independent (prev, cur) pairs; stream i's content moves by shifts[i] = (dx, dy) px.

    Returns prev[n,h,w], cur[n,h,w] (uint8) and shifts[n,2] (float32; x.5 entries mark
    half-pixel pairs made by a truncating average of two integer crops).
    """
    rng = np.random.default_rng(seed)
    m = max_shift + 1
    tex = box_blur3(rng.integers(0, 256, size=(n, h + 2 * m, w + 2 * m), dtype=np.uint8))
    sh = rng.integers(-max_shift, max_shift + 1, size=(n, 2))
    prev = np.ascontiguousarray(tex[:, m:m + h, m:m + w])
    cur = np.empty_like(prev)
    shifts = sh.astype(np.float32)
    for i in range(n):
        dx, dy = int(sh[i, 0]), int(sh[i, 1])
        c = tex[i, m - dy:m - dy + h, m - dx:m - dx + w]
        if half_pixel_every and i % half_pixel_every == 0 and dx < max_shift:
            c2 = tex[i, m - dy:m - dy + h, m - dx - 1:m - dx - 1 + w]
            c = ((c.astype(np.uint16) + c2 + 1) >> 1).astype(np.uint8)
            shifts[i, 0] += 0.5
        cur[i] = c
    return prev, cur, shifts


def gaussian_texture(h, w, seed=0, sigma=2.0):
    """Seeded noise low-passed with a separable Gaussian, re-normalised to 0..255 (LK frames)."""
    rng = np.random.default_rng(seed)
    a = rng.random((h, w)).astype(np.float64)
    r = int(3 * sigma + 0.5)
    k = np.exp(-0.5 * (np.arange(-r, r + 1) / sigma) ** 2)
    k /= k.sum()
    a = np.apply_along_axis(lambda v: np.convolve(np.pad(v, r, mode="wrap"), k, mode="valid"), 1, a)
    a = np.apply_along_axis(lambda v: np.convolve(np.pad(v, r, mode="wrap"), k, mode="valid"), 0, a)
    a -= a.min()
    a /= a.max()
    return np.round(a * 255.0).astype(np.uint8)


def lk_sequence(n_frames, w=640, h=480, dx=-3, dy=2, seed=0, sigma=2.0):
    """Frames of one stream: content moves (+dx,+dy) px per frame."""
    m = (abs(dx) + abs(dy)) * n_frames + 8
    tex = gaussian_texture(h + 2 * m, w + 2 * m, seed, sigma)
    return [np.ascontiguousarray(tex[m - k * dy:m - k * dy + h, m - k * dx:m - k * dx + w]) for k in range(n_frames)]
