"""Seeded synthetic inputs (numpy only, so they are identical here and on the GPU box).

Frames follow SURVEY.md 8(d): quarter-resolution uniform noise, nearest-upsampled x4,
smoothed with a 5-tap binomial filter, plus random filled rectangles and lines. That gives
corner-rich but not saturated images (pure full-res noise is unrepresentative).
"""
import numpy as np


def _smooth(img):
    k = np.array([1, 4, 6, 4, 1], dtype=np.uint32)
    p = np.pad(img.astype(np.uint32), 2, mode="reflect")
    h = sum(k[i] * p[:, i:i + img.shape[1]] for i in range(5))
    v = sum(k[i] * h[i:i + img.shape[0], :] for i in range(5))
    return ((v + 128) >> 8).astype(np.uint8)


def frame(seed, w=640, h=480, kind="texture"):
    """Return an (h, w) uint8 image."""
    rng = np.random.default_rng(seed)
    if kind == "black":
        return np.zeros((h, w), np.uint8)
    if kind == "gradient":  # weak corners only: forces the minThFAST retry
        yy, xx = np.mgrid[0:h, 0:w]
        img = ((xx * 3 + yy * 2) // 8 % 256).astype(np.uint8)
        n = rng.integers(0, 18, size=(h // 8 + 1, w // 8 + 1)).astype(np.uint8)
        img = img + np.kron(n, np.ones((8, 8), np.uint8))[:h, :w]
        return img.astype(np.uint8)
    if kind == "noise":
        return rng.integers(0, 256, size=(h, w), dtype=np.uint8)
    q = rng.integers(0, 256, size=((h + 3) // 4, (w + 3) // 4), dtype=np.uint8)
    img = np.kron(q, np.ones((4, 4), np.uint8))[:h, :w]
    img = _smooth(img)
    for _ in range(50):
        x0, y0 = int(rng.integers(0, w - 8)), int(rng.integers(0, h - 8))
        rw, rh = int(rng.integers(4, max(5, w // 6))), int(rng.integers(4, max(5, h // 6)))
        val = int(rng.integers(0, 256))
        if rng.random() < 0.7:
            img[y0:y0 + rh, x0:x0 + rw] = val
        else:  # thick diagonal line
            n = min(rw, rh)
            for t in range(n):
                img[min(h - 1, y0 + t):min(h, y0 + t + 3), min(w - 1, x0 + t):min(w, x0 + t + 3)] = val
    return np.ascontiguousarray(img)


def descriptors(seed, n, flip_p=None, base=None):
    """Random 256-bit descriptors (n,32) uint8; with base/flip_p: perturbed copies of base."""
    rng = np.random.default_rng(seed)
    if base is None:
        return rng.integers(0, 256, size=(n, 32), dtype=np.uint8)
    bits = np.unpackbits(base[:n], axis=1)
    flips = (rng.random(bits.shape) < flip_p).astype(np.uint8)
    return np.packbits(bits ^ flips, axis=1)
