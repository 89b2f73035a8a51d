"""Deterministic synthetic inputs for tests and bench.

Images: mid-grey canvas, multi-scale random rectangles, low uniform noise — corner-rich enough that every pyramid level
exceeds its keypoint quota, so the quadtree's second phase and its tie-break are exercised.
Descriptors: uniform random 256-bit rows; queries are database rows with Binomial(256, 0.1) flipped bits.
"""
import numpy as np


def synth_image(h, w, seed):
    rng = np.random.default_rng(seed)
    img = np.full((h, w), 128, np.uint8)
    for s in (64, 32, 16, 8):
        n = (h * w) // (2 * s * s)
        x0 = rng.integers(0, w, n)
        y0 = rng.integers(0, h, n)
        ws = rng.integers(s // 2, 2 * s, n)
        hs = rng.integers(s // 2, 2 * s, n)
        g = rng.integers(0, 256, n)
        for i in range(n):
            img[y0[i]:y0[i] + hs[i], x0[i]:x0[i] + ws[i]] = g[i]
    noise = rng.integers(0, 6, (h, w))
    return np.minimum(img.astype(np.int32) + noise, 255).astype(np.uint8)


def noise_image(h, w, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w), dtype=np.uint8)


def synth_stereo_pair(h, w, seed):
    """Left image + right image = left shifted by a per-row-band constant integer disparity (4..64 px)."""
    left = synth_image(h, w, seed)
    rng = np.random.default_rng(seed + 7919)
    right = np.empty_like(left)
    band = 48
    for y0 in range(0, h, band):
        d = int(rng.integers(4, 65))
        rows = slice(y0, min(y0 + band, h))
        right[rows, :w - d] = left[rows, d:]
        right[rows, w - d:] = left[rows, w - d - 1:w - d]
    return left, right


def synth_descriptors(n_db, n_q, seed=5, flip_p=0.1):
    rng = np.random.default_rng(seed)
    db = rng.integers(0, 256, (n_db, 32), dtype=np.uint8)
    src = rng.integers(0, n_db, n_q)
    flips = rng.random((n_q, 256)) < flip_p
    q = db[src] ^ np.packbits(flips, axis=1, bitorder="little")
    return db, q, src
