"""Synthetic inputs for parity tests and benchmarks (no datasets are available offline).

Definitions follow SURVEY.md §8(d): seeded multi-octave noise + random filled rectangles/discs for strong corners +
one low-contrast quadrant so the minThFAST retry path fires; stereo right image = left warped by a smooth horizontal
disparity field + independent noise; descriptor sets = random bytes with planted near- and exact duplicates.
Pure numpy, deterministic per seed.
"""
import numpy as np


def _upsample_bilinear(g, h, w):
    gh, gw = g.shape
    ys = np.linspace(0, gh - 1, h)
    xs = np.linspace(0, gw - 1, w)
    y0 = np.floor(ys).astype(int); x0 = np.floor(xs).astype(int)
    y1 = np.minimum(y0 + 1, gh - 1); x1 = np.minimum(x0 + 1, gw - 1)
    fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
    a = g[y0][:, x0]; b = g[y0][:, x1]; c = g[y1][:, x0]; d = g[y1][:, x1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def synth_image(w, h, seed=0):
    """One textured 8-bit frame of the given shape."""
    rng = np.random.default_rng(seed)
    img = np.zeros((h, w), np.float64)
    for cell, wt in ((8, 1.0), (16, 0.5), (32, 0.25), (64, 0.125)):
        g = rng.random((h // cell + 2, w // cell + 2))
        img += wt * _upsample_bilinear(g, h, w)
    img = (img - img.mean()) / (img.std() + 1e-9) * 45.0 + 128.0
    nshape = int(200 * (w * h / 307200.0))
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(nshape):
        s = int(rng.integers(6, 41)); cx = int(rng.integers(0, w)); cy = int(rng.integers(0, h))
        v = float(rng.integers(0, 256))
        if rng.random() < 0.5:
            img[max(cy - s // 2, 0):cy + s // 2, max(cx - s // 2, 0):cx + s // 2] = v
        else:
            y0, y1 = max(cy - s, 0), min(cy + s + 1, h); x0, x1 = max(cx - s, 0), min(cx + s + 1, w)
            m = (yy[y0:y1, x0:x1] - cy) ** 2 + (xx[y0:y1, x0:x1] - cx) ** 2 <= (s // 2) ** 2
            img[y0:y1, x0:x1][m] = v
    # low-contrast quadrant (bottom-right): contrast x0.15 around 128
    q = img[h // 2:, w // 2:]
    img[h // 2:, w // 2:] = (q - 128.0) * 0.15 + 128.0
    img += rng.normal(0, 1.5, img.shape)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def synth_batch(w, h, n, seed0=0):
    return np.stack([synth_image(w, h, seed0 + i) for i in range(n)])


def synth_stereo_pair(w, h, seed=0, dmin=2.0, dmax=80.0):
    """(left, right): right(x, y) = left(x + d(x, y), y) with a smooth disparity field, + noise sigma 2."""
    left = synth_image(w, h, seed)
    rng = np.random.default_rng(seed + 1000)
    g = rng.random((h // 64 + 2, w // 64 + 2))
    d = dmin + (dmax - dmin) * _upsample_bilinear(g, h, w)
    xs = np.arange(w)[None, :] + d
    x0 = np.clip(np.floor(xs).astype(int), 0, w - 1); x1 = np.clip(x0 + 1, 0, w - 1)
    fx = xs - np.floor(xs)
    rows = np.arange(h)[:, None]
    L = left.astype(np.float64)
    right = L[rows, x0] * (1 - fx) + L[rows, x1] * fx + rng.normal(0, 2.0, (h, w))
    return left, np.clip(np.rint(right), 0, 255).astype(np.uint8)


def synth_descriptors(n_train, n_query, seed=7, dup=1000, flip_p=0.08, matched_frac=0.6):
    """Train set of random 256-bit descriptors with `dup` planted exact duplicates; queries: `matched_frac` are a
    train row with Binomial(256, flip_p) bits flipped, the rest random."""
    rng = np.random.default_rng(seed)
    train = rng.integers(0, 256, (n_train, 32), dtype=np.uint8)
    if dup and n_train > 2 * dup:
        src = rng.integers(0, n_train, dup); dst = rng.integers(0, n_train, dup)
        train[dst] = train[src]
    q = rng.integers(0, 256, (n_query, 32), dtype=np.uint8)
    nm = int(n_query * matched_frac)
    rows = rng.integers(0, n_train, nm)
    flips = (rng.random((nm, 256)) < flip_p)
    q[:nm] = train[rows] ^ np.packbits(flips, axis=1, bitorder="little")
    perm = rng.permutation(n_query)
    return train, np.ascontiguousarray(q[perm])
