"""Deterministic synthetic inputs for the ORB front end.

Frames: mid-grey canvas, random filled rotated rectangles, 3x3 Gaussian smoothing,
additive Gaussian noise, clipped to u8.  Pure numpy (PCG64 seeds) so the same bytes are
produced in the build container and on the GPU box.  Not part of the hot path.
"""
from __future__ import annotations

import numpy as np


def _blur3(img: np.ndarray) -> np.ndarray:
    # separable 3-tap Gaussian, sigma 0.8, reflect-101 border, float32
    k = np.exp(-0.5 * (np.arange(-1, 2) / 0.8) ** 2).astype(np.float32)
    k /= k.sum()
    p = np.pad(img, 1, mode="reflect")
    h = k[0] * p[:, :-2] + k[1] * p[:, 1:-1] + k[2] * p[:, 2:]
    return k[0] * h[:-2] + k[1] * h[1:-1] + k[2] * h[2:]


def canvas(width: int, height: int, seed: int, nrect: int | None = None) -> np.ndarray:
    """Noise-free float32 canvas of filled rotated rectangles (values 0..255)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    if nrect is None:
        nrect = max(1, int(round(400.0 * width * height / (640.0 * 480.0))))
    img = np.full((height, width), 128.0, np.float32)
    cx = rng.uniform(0, width, nrect)
    cy = rng.uniform(0, height, nrect)
    sw = rng.uniform(8, 120, nrect)
    sh = rng.uniform(8, 120, nrect)
    ang = rng.uniform(0, np.pi, nrect)
    grey = rng.integers(0, 256, nrect)
    for i in range(nrect):
        c, s = np.cos(ang[i]), np.sin(ang[i])
        hw, hh = sw[i] / 2, sh[i] / 2
        rad = int(np.ceil(np.hypot(hw, hh))) + 1
        x0, x1 = max(0, int(cx[i]) - rad), min(width, int(cx[i]) + rad + 1)
        y0, y1 = max(0, int(cy[i]) - rad), min(height, int(cy[i]) + rad + 1)
        if x0 >= x1 or y0 >= y1:
            continue
        xs = np.arange(x0, x1, dtype=np.float32) - np.float32(round(cx[i]))
        ys = np.arange(y0, y1, dtype=np.float32) - np.float32(round(cy[i]))
        X, Y = np.meshgrid(xs, ys)
        u = X * c + Y * s
        v = -X * s + Y * c
        m = (np.abs(u) <= hw) & (np.abs(v) <= hh)
        img[y0:y1, x0:x1][m] = grey[i]
    return _blur3(img)


def frame(width: int = 640, height: int = 480, seed: int = 0) -> np.ndarray:
    """One synthetic u8 frame (config 1 uses seed 0 at 640x480)."""
    rng = np.random.Generator(np.random.PCG64(seed + 0x5EED))
    img = canvas(width, height, seed)
    img = img + rng.normal(0.0, 2.0, img.shape).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def sequence(n: int, width: int = 640, height: int = 480, seed: int = 0, max_shift: int = 8) -> tuple[np.ndarray, np.ndarray]:
    """Monocular stream (config 2): n frames cropped from one large canvas with a random-walk
    translation of at most +-max_shift px per frame, independent noise per frame.
    Returns (frames[n,h,w] u8, offsets[n,2] int (x,y) of each crop)."""
    rng = np.random.Generator(np.random.PCG64(seed + 0xABCD))
    margin = 256
    cw, ch = width + 2 * margin, height + 2 * margin
    base = canvas(cw, ch, seed)
    pos = np.zeros((n, 2), np.int64)
    p = np.array([margin, margin], np.int64)
    for i in range(n):
        pos[i] = p
        step = rng.integers(-max_shift, max_shift + 1, 2)
        q = p + step
        q[0] = min(max(q[0], 0), cw - width)
        q[1] = min(max(q[1], 0), ch - height)
        p = q
    out = np.empty((n, height, width), np.uint8)
    for i in range(n):
        x, y = pos[i]
        crop = base[y:y + height, x:x + width]
        noise = rng.normal(0.0, 2.0, crop.shape).astype(np.float32)
        out[i] = np.clip(np.rint(crop + noise), 0, 255).astype(np.uint8)
    return out, pos


def stereo_pair(width: int = 1241, height: int = 376, seed: int = 0, disparity: int = 20) -> tuple[np.ndarray, np.ndarray]:
    """Stereo pair (config 3): one scene seen with a global disparity, uL - uR = disparity > 0 (a scene point sits
    further left in the right image, as for a rectified pair)."""
    rng = np.random.Generator(np.random.PCG64(seed + 0x57E0))
    base = canvas(width + disparity, height, seed)
    left = base[:, :width]
    right = base[:, disparity:disparity + width]
    nl = rng.normal(0.0, 2.0, left.shape).astype(np.float32)
    nr = rng.normal(0.0, 2.0, left.shape).astype(np.float32)
    return (np.clip(np.rint(left + nl), 0, 255).astype(np.uint8),
            np.clip(np.rint(right + nr), 0, 255).astype(np.uint8))


def noise_frame(width: int = 640, height: int = 480, seed: int = 0) -> np.ndarray:
    """Uniform-noise frame: saturates every FAST cell (degenerate fixture)."""
    rng = np.random.Generator(np.random.PCG64(seed + 0x4015E))
    return rng.integers(0, 256, (height, width), dtype=np.uint8)


def descriptors(n: int, seed: int) -> np.ndarray:
    """n x 32 u8 i.i.d. uniform descriptor rows."""
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.integers(0, 256, (n, 32), dtype=np.uint8)


def planted_descriptors(train: np.ndarray, m: int, seed: int, frac: float = 0.1, max_flips: int = 30) -> np.ndarray:
    """m query rows: i.i.d. uniform, with a fraction replaced by near-duplicates of random train
    rows (<= max_flips flipped bits) so the ratio test has positives (config 5)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    q = rng.integers(0, 256, (m, 32), dtype=np.uint8)
    k = int(m * frac)
    rows = rng.choice(m, k, replace=False)
    src = rng.integers(0, train.shape[0], k)
    dup = train[src].copy()
    nflip = rng.integers(0, max_flips + 1, k)
    for i in range(k):
        bits = rng.choice(256, nflip[i], replace=False)
        for b in bits:
            dup[i, b >> 3] ^= np.uint8(1 << (b & 7))
    q[rows] = dup
    return q


def planted_descriptors_fast(train: np.ndarray, m: int, seed: int, frac: float = 0.1, max_flips: int = 30):
    """Vectorised planted_descriptors for the 1M x 1M shape of config 5.  Returns (query, rows, src, nflip): query
    rows `rows` are train rows `src` with exactly `nflip` distinct bits flipped; all other rows are i.i.d. uniform."""
    rng = np.random.Generator(np.random.PCG64(seed))
    q = rng.integers(0, 256, (m, 32), dtype=np.uint8)
    k = int(m * frac)
    rows = np.sort(rng.choice(m, k, replace=False))
    src = rng.integers(0, train.shape[0], k)
    nflip = rng.integers(0, max_flips + 1, k)
    dup = train[src].copy()
    for c0 in range(0, k, 8192):                      # distinct bit positions per row: the smallest keys of a random table
        c1 = min(k, c0 + 8192)
        pos = np.argsort(rng.random((c1 - c0, 256)), axis=1)[:, :max_flips]
        use = np.arange(max_flips)[None, :] < nflip[c0:c1, None]
        bits = np.zeros((c1 - c0, 256), np.uint8)
        r = np.repeat(np.arange(c1 - c0), max_flips)[use.ravel()]
        bits[r, pos.ravel()[use.ravel()]] = 1
        dup[c0:c1] ^= np.packbits(bits, axis=1, bitorder="little")
    q[rows] = dup
    return q, rows, src, nflip
