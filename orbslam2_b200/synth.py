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


def synth_tracking_pair(n_points=1200, seed=0, width=640, height=480, stereo=True, motion=(0.05, -0.02, 0.12),
                        conflicts=150, nlevels=8, scale=1.2):
    """A (LastFrame, CurrentFrame) pair for ORBmatcher::SearchByProjection(Frame, Frame): random 3-D points seen by two
    nearby camera poses. Current keypoints = projections + pixel noise + distractors (shuffled, so grid order differs from
    query order); descriptors = the map point's descriptor with a few bits flipped. `conflicts` last-frame points are
    near-duplicates of others (same place, similar descriptor) so that several queries compete for one keypoint and
    the greedy claim order matters; ~15 % of the map points have no observations (temporal points that do not block)."""
    from .frame import Frame
    from .extractor import KP_DTYPE
    rng = np.random.default_rng(seed)
    fx = fy = np.float32(520.0); cx, cy = np.float32(width / 2 - 0.5), np.float32(height / 2 + 0.25)
    bf = np.float32(40.0)
    sf = np.ones(nlevels, np.float32)
    for i in range(1, nlevels):
        sf[i] = np.float32(np.float64(sf[i - 1]) * np.float64(np.float32(scale)))
    # world points in front of the last camera (identity pose)
    z = rng.uniform(2.0, 12.0, n_points).astype(np.float32)
    u = rng.uniform(20, width - 20, n_points).astype(np.float32); v = rng.uniform(20, height - 20, n_points).astype(np.float32)
    X = np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], 1).astype(np.float32)
    desc = rng.integers(0, 256, (n_points, 32), dtype=np.uint8)
    octave = rng.integers(0, nlevels, n_points).astype(np.int32)
    angle = rng.uniform(0, 360, n_points).astype(np.float32)
    # near-duplicate map points
    dup = rng.choice(n_points, conflicts, replace=False)
    src = rng.choice(np.setdiff1d(np.arange(n_points), dup), conflicts)
    X[dup] = X[src] + rng.normal(0, 0.002, (conflicts, 3)).astype(np.float32)
    desc[dup] = desc[src]
    flip = rng.integers(0, 256, (conflicts, 2))
    for k, d in enumerate(dup):
        for b in flip[k]:
            desc[d, b >> 3] ^= np.uint8(1 << (b & 7))
    octave[dup] = octave[src]; angle[dup] = angle[src] + rng.normal(0, 2, conflicts).astype(np.float32)
    angle = np.mod(angle, 360).astype(np.float32)

    def kps(x, y, o, a):
        k = np.zeros(len(x), KP_DTYPE)
        k["x"], k["y"], k["octave"], k["angle"] = x, y, o, a
        k["size"] = 31 * sf[o]; k["class_id"] = -1
        return k

    last = Frame(kps(u, v, octave, angle), desc, sf, fx, fy, cx, cy, bf, width, height,
                 u_right=(u - bf / z).astype(np.float32) if stereo else None)
    last.mp_valid[:] = 1
    last.mp_valid[rng.random(n_points) < 0.05] = 0
    last.mvbOutlier[rng.random(n_points) < 0.03] = 1
    last.mp_observed[:] = (rng.random(n_points) > 0.15).astype(np.uint8)
    last.mp_world[:] = X
    last.mp_desc[:] = desc
    # current pose: small rotation about y + translation
    th = 0.02
    R = np.array([[np.cos(th), 0, np.sin(th)], [0, 1, 0], [-np.sin(th), 0, np.cos(th)]], np.float32)
    T = np.eye(4, dtype=np.float32); T[:3, :3] = R; T[:3, 3] = np.asarray(motion, np.float32)
    Xc = (R.astype(np.float64) @ X.T.astype(np.float64)).T + np.asarray(motion, np.float64)
    uc = (fx * Xc[:, 0] / Xc[:, 2] + cx + rng.normal(0, 1.5, n_points)).astype(np.float32)
    vc = (fy * Xc[:, 1] / Xc[:, 2] + cy + rng.normal(0, 1.5, n_points)).astype(np.float32)
    keep = np.setdiff1d(np.arange(n_points), dup)              # one keypoint per physical point
    keep = keep[rng.random(len(keep)) > 0.1]                   # some points are not re-detected
    nd = n_points // 3                                         # distractors
    cu = np.concatenate([uc[keep], rng.uniform(0, width, nd).astype(np.float32)])
    cv = np.concatenate([vc[keep], rng.uniform(0, height, nd).astype(np.float32)])
    cd = np.concatenate([desc[keep], rng.integers(0, 256, (nd, 32), dtype=np.uint8)])
    nbits = rng.binomial(256, 0.06, len(keep))
    for k in range(len(keep)):
        for b in rng.integers(0, 256, nbits[k]):
            cd[k, b >> 3] ^= np.uint8(1 << (b & 7))
    co = np.concatenate([np.clip(octave[keep] + rng.integers(-1, 2, len(keep)), 0, nlevels - 1),
                         rng.integers(0, nlevels, nd)]).astype(np.int32)
    ca = np.concatenate([np.mod(angle[keep] + rng.normal(0, 4, len(keep)), 360),
                         rng.uniform(0, 360, nd)]).astype(np.float32)
    wild = rng.random(len(keep)) < 0.08                        # matches the rotation-consistency filter must reject
    ca[:len(keep)][wild] = rng.uniform(0, 360, int(wild.sum())).astype(np.float32)
    cz = np.concatenate([Xc[keep, 2], rng.uniform(2, 12, nd)])
    cur_ur = (cu - bf / cz).astype(np.float32)
    cur_ur[rng.random(len(cu)) < 0.2] = -1                     # keypoints without a stereo match
    perm = rng.permutation(len(cu))
    cur = Frame(kps(cu[perm], cv[perm], co[perm], ca[perm]), cd[perm], sf, fx, fy, cx, cy, bf, width, height,
                u_right=cur_ur[perm] if stereo else None)
    cur.SetPose(T)
    return last, cur


FBOW_SIG = 55824124          # Thirdparty/fbow/src/fbow.cpp:166
FBOW_PARAMS_BYTES = 120      # sizeof(fbow::Vocabulary::params), Thirdparty/fbow/src/fbow.h:115-125 (x86-64 layout)


def _roundup(x, a):
    return (x + a - 1) // a * a


def synth_vocabulary(k=10, levels=3, seed=0, align=8, ragged=False, leaf_prob=0.0, name="orb"):
    """Bytes of a synthetic vocabulary in the reference's `.fbow` wire format (Vocabulary::toStream,
    Thirdparty/fbow/src/fbow.cpp:164-171; block layout fbow.h:160-190; sizes as setParams computes them,
    fbow.cpp:10-49). `Vocabulary/orb_mur.fbow` itself is not in the reference tree, so tests and benches use this.

    A k-ary tree of `levels` levels: a node's descriptor is its parent's with a level-dependent number of random bit
    flips (so the descent is a meaningful nearest-centre search). ragged: blocks hold 1..k nodes; leaf_prob: inner
    nodes that are leaves before the last level (words at mixed depths). Word ids run in creation order; weights are
    random positive floats."""
    rng = np.random.default_rng(seed)
    desc_wp = _roundup(32, align)
    feat_off = _roundup(8, align)
    child_off = feat_off + k * desc_wp
    block_size = _roundup(feat_off + k * (desc_wp + 8), align)
    parents = rng.integers(0, 256, (1, 32), dtype=np.uint8)   # the (virtual) root descriptor
    parent_block = np.zeros(1, np.uint32)
    base, word = 0, 0
    chunks = []
    for d in range(levels):
        nb = len(parents)
        n = rng.integers(1, k + 1, nb) if ragged else np.full(nb, k)
        if d == 0:
            n[0] = k
        mask = rng.integers(0, 256, (nb, k, 32), dtype=np.uint8)
        for _ in range(d + 1):                                  # bit density 2^-(d+2)
            mask &= rng.integers(0, 256, (nb, k, 32), dtype=np.uint8)
        desc = parents[:, None, :] ^ mask
        valid = np.arange(k)[None, :] < n[:, None]
        leaf = np.ones((nb, k), bool) if d == levels - 1 else rng.random((nb, k)) < leaf_prob
        inner = valid & ~leaf
        leafv = valid & leaf
        info = np.zeros((nb, k, 2), np.uint32)
        n_inner, n_leaf = int(inner.sum()), int(leafv.sum())
        info[..., 0][inner] = base + nb + np.arange(n_inner, dtype=np.uint32)
        info[..., 0][leafv] = (word + np.arange(n_leaf, dtype=np.uint32)) | np.uint32(0x80000000)
        info[..., 1][leafv] = (0.25 + 8.0 * rng.random(n_leaf)).astype(np.float32).view(np.uint32)
        blk = np.zeros((nb, block_size), np.uint8)
        blk[:, 0:2] = n.astype(np.uint16)[:, None].view(np.uint8)
        blk[:, 2:4] = (~inner.any(axis=1)).astype(np.uint16)[:, None].view(np.uint8)
        blk[:, 4:8] = parent_block.astype(np.uint32)[:, None].view(np.uint8)
        feats = np.zeros((nb, k, desc_wp), np.uint8)
        feats[:, :, :32] = desc * valid[:, :, None]
        blk[:, feat_off:child_off] = feats.reshape(nb, k * desc_wp)
        blk[:, child_off:child_off + 8 * k] = info.reshape(nb, 2 * k).view(np.uint8)
        chunks.append(blk)
        bi, _ = np.nonzero(inner)
        parents = desc[inner]
        parent_block = (base + bi).astype(np.uint32)
        base += nb
        word += n_leaf
        if n_inner == 0:
            break
    data = np.concatenate(chunks)
    nblocks = len(data)
    hdr = np.zeros(FBOW_PARAMS_BYTES, np.uint8)
    nm = name.encode()[:49]
    hdr[:len(nm)] = np.frombuffer(nm, np.uint8)
    hdr[52:60] = np.array([align, nblocks], np.uint32).view(np.uint8)
    hdr[64:104] = np.array([desc_wp, block_size, feat_off, child_off, nblocks * block_size], np.uint64).view(np.uint8)
    hdr[104:112] = np.array([0, 32], np.int32).view(np.uint8)   # CV_8UC1, 32 bytes
    hdr[112:116] = np.array([k], np.uint32).view(np.uint8)
    return np.array([FBOW_SIG], np.uint64).tobytes() + hdr.tobytes() + data.tobytes()


def synth_vocabulary_features(blob, n, seed=0, flips=12, random_frac=0.1):
    """n descriptors that resemble a vocabulary's words: a random leaf descriptor with `flips` random bits toggled
    (a fraction random_frac is uniform noise). Reads the wire format only (layout as in synth_vocabulary)."""
    rng = np.random.default_rng(seed)
    raw = np.frombuffer(blob, np.uint8)
    hdr = raw[8:8 + FBOW_PARAMS_BYTES]
    nblocks = int(hdr[56:60].view(np.uint32)[0])
    desc_wp, block_size, feat_off, child_off = (int(v) for v in hdr[64:96].view(np.uint64))
    k = int(hdr[112:116].view(np.uint32)[0])
    data = raw[8 + FBOW_PARAMS_BYTES:8 + FBOW_PARAMS_BYTES + nblocks * block_size].reshape(nblocks, block_size)
    cnt = data[:, 0:2].copy().view(np.uint16)[:, 0]
    info = data[:, child_off:child_off + 8 * k].copy().view(np.uint32).reshape(nblocks, k, 2)
    leaf = ((info[..., 0] & np.uint32(0x80000000)) != 0) & (np.arange(k)[None, :] < cnt[:, None])
    bi, ci = np.nonzero(leaf)
    pick = rng.integers(0, len(bi), n)
    feats = data[:, feat_off:feat_off + k * desc_wp].reshape(nblocks, k, desc_wp)[bi[pick], ci[pick], :32].copy()
    bits = np.unpackbits(feats, axis=1)
    for _ in range(flips):
        bits[np.arange(n), rng.integers(0, 256, n)] ^= 1
    feats = np.packbits(bits, axis=1)
    noise = rng.random(n) < random_frac
    feats[noise] = rng.integers(0, 256, (int(noise.sum()), 32), dtype=np.uint8)
    return feats
