"""CPU tests that PIN THE ORACLE (no GPU needed).

1. Every OpenCV primitive restated in oracle/cvprim.hpp is checked bit-for-bit against known-answer vectors generated
   from the real OpenCV (cv2 4.13.0) by tests/golden/make_golden.py -> tests/golden/cv2_kat.npz.
2. The whole-extractor restatement (oracle/orb_oracle.cpp) is checked against the UNMODIFIED reference
   /root/reference/src/ORBextractor.cc compiled on the cv shim (oracle/_ref): pyramids identical; keypoints identical
   except for the documented allocator-dependent quadtree tie-break, and on every keypoint both select,
   angle / response / size / descriptor are identical.
3. Hand-derived micro cases (DescriptorDistance, scan tie rules, table values quoted in SURVEY.md §8 a1).
"""
import os

import numpy as np
import pytest

from oracle_lib import Oracle, OraclePrims, RefExtractor, canonical, have_ref, TUM1
from orbslam2_b200.synth import synth_image, synth_descriptors

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cv2_kat.npz")
NAMES = ["synth", "noise", "steps"]


@pytest.fixture(scope="module")
def kat():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def P():
    return OraclePrims()


@pytest.mark.parametrize("name", NAMES)
def test_resize_matches_cv2(kat, P, name):
    prev = kat[f"img_{name}"]
    for l in range(1, 4):
        exp = kat[f"resize_{name}_{l}"]
        got = P.resize(prev, exp.shape[1], exp.shape[0])
        assert np.array_equal(got, exp), f"{name} level {l}: {(got != exp).sum()} px differ"
        prev = exp
    exp = kat[f"resize_{name}_odd"]
    assert np.array_equal(P.resize(kat[f"img_{name}"], exp.shape[1], exp.shape[0]), exp)


@pytest.mark.parametrize("name", NAMES)
def test_border_and_blur_match_cv2(kat, P, name):
    img = kat[f"img_{name}"]
    assert np.array_equal(P.border(img, 19), kat[f"border_{name}"])
    assert np.array_equal(P.blur(img), kat[f"blur_{name}"])


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("th", [7, 20])
def test_fast_matches_cv2(kat, P, name, th):
    """Set, score AND row-major order of cv::FAST(th, nonmax=true) TYPE_9_16."""
    exp = kat[f"fast_{name}_{th}"]
    got = P.fast(kat[f"img_{name}"], th)
    assert np.array_equal(got, exp), f"{len(got)} vs {len(exp)} keypoints"


def test_fast_atan2_matches_cv2(kat, P):
    got = P.fast_atan2(kat["atan2_y"], kat["atan2_x"])
    assert np.array_equal(got, kat["atan2"])  # bit-exact fp32


def test_golden_was_made_with_opencv_4_13(kat):
    assert str(kat["cv2_version"]).startswith("4.")


# ---------------------------------------------------------------------------------------------- oracle vs reference
needs_ref = pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (reference tree absent)")


def _compare_with_reference(orc, ref, img, cap, nlevels):
    """Per level: (identical?, tie at the cut?, #oracle, #reference). Common keypoints must agree in every field."""
    ko, do = canonical(*orc.extract(img, cap))
    kr, dr = canonical(*ref.extract(img, cap))
    out = []
    for l in range(nlevels):
        a, b = ko[ko["octave"] == l], kr[kr["octave"] == l]
        da, db = do[ko["octave"] == l], dr[kr["octave"] == l]
        same = len(a) == len(b) and a.tobytes() == b.tobytes() and np.array_equal(da, db)
        out.append((same, orc.tie_at_cut(l), len(a), len(b)))
    key = lambda k: (int(k["octave"]), float(k["x"]), float(k["y"]))  # noqa: E731
    io = {key(k): i for i, k in enumerate(ko)}
    ir = {key(k): i for i, k in enumerate(kr)}
    common = sorted(set(io) & set(ir))
    for c in common:
        assert ko[io[c]].tobytes() == kr[ir[c]].tobytes(), f"keypoint fields differ at {c}"
        assert np.array_equal(do[io[c]], dr[ir[c]]), f"descriptor differs at {c}"
    return out, len(common), len(kr)


@needs_ref
@pytest.mark.parametrize("w,h,nf,seeds", [(640, 480, 1000, 50), (752, 480, 1200, 12), (1241, 376, 2000, 12), (208, 160, 300, 50),
                                          (376, 241, 500, 30), (480, 640, 1000, 12), (641, 479, 1000, 8)])
def test_oracle_vs_reference_extractor(w, h, nf, seeds):
    """Oracle vs the UNMODIFIED src/ORBextractor.cc: tables and padded pyramids identical; on every level whose last
    sorted quadtree sweep is NOT cut between two nodes of equal key count, the selected keypoints and descriptors are
    identical byte for byte. Where the cut does fall inside such a tie the reference's choice depends on heap addresses
    (std::sort on pair<int, ExtractorNode*>, src/ORBextractor.cc:678); the oracle breaks the tie by (y, x) — SURVEY.md §7
    H1 — and only those levels may differ, by a handful of keypoints."""
    p = dict(TUM1); p["nfeatures"] = nf
    orc, ref = Oracle(**p), RefExtractor(**p)
    t_o, pl_o, um_o = orc.tables(); t_r, pl_r, um_r = ref.tables()
    assert np.array_equal(t_o, t_r) and np.array_equal(pl_o, pl_r) and np.array_equal(um_o, um_r)
    levels = ties = exact_tied = 0
    common_total = ref_total = 0
    for seed in range(seeds):
        img = synth_image(w, h, seed)
        res, common, nref = _compare_with_reference(orc, ref, img, 4 * nf, p["nlevels"])
        if seed < 2:
            for l in range(p["nlevels"]):
                assert orc.level_size(l) == ref.level_size(l)
                assert np.array_equal(orc.level(l, True), ref.level(l, True)), f"padded pyramid level {l}"
        for l, (same, tie, no, nr) in enumerate(res):
            levels += 1; ties += tie
            if not tie:
                assert same, f"seed {seed} level {l}: no tie at the quadtree cut, yet oracle ({no}) != reference ({nr})"
            else:
                exact_tied += same
                assert abs(no - nr) <= 3, f"seed {seed} level {l}: counts {no} vs {nr}"
        common_total += common; ref_total += nref
    assert common_total >= 0.97 * ref_total
    print(f"{w}x{h}: {ties}/{levels} levels with a tie at the cut ({exact_tied} of them identical anyway); "
          f"{common_total}/{ref_total} keypoints in common")


@needs_ref
def test_oracle_vs_reference_extractor_4k():
    p = dict(TUM1); p["nfeatures"] = 8000
    orc, ref = Oracle(**p), RefExtractor(**p)
    res, common, nref = _compare_with_reference(orc, ref, synth_image(3840, 2160, 0), 32000, 8)
    for l, (same, tie, no, nr) in enumerate(res):
        assert same or tie, f"4K level {l}: no tie at the cut, yet oracle ({no}) != reference ({nr})"
    assert common >= 0.97 * nref


@needs_ref
def test_reference_order_is_level_major():
    ref = RefExtractor(**TUM1)
    k, _ = ref.extract(synth_image(640, 480, 2), 4000)
    assert np.all(np.diff(k["octave"]) >= 0)
    assert set(np.unique(k["class_id"])) == {-1}


# ---------------------------------------------------------------------------------------------- micro cases
def test_tables_quoted_in_survey():
    orc = Oracle(**TUM1)
    t, pl, um = orc.tables()
    assert list(pl) == [217, 181, 151, 126, 105, 87, 73, 60]                      # SURVEY.md §8 geometry table
    assert list(um) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert t[0][0] == 1.0 and abs(t[0][1] - 1.2) < 1e-6
    sizes = [orc_level for orc_level in [(640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193),
                                         (214, 161), (179, 134)]]
    orc.extract(synth_image(640, 480, 0), 4000)
    assert [orc.level_size(l) for l in range(8)] == sizes


def test_descriptor_distance_micro(P):
    z = np.zeros(32, np.uint8); o = np.full(32, 255, np.uint8)
    assert P.hamming(z, o) == 256 and P.hamming(z, z) == 0
    a = z.copy(); a[5] = 0b10010001
    assert P.hamming(a, z) == 3
    rng = np.random.default_rng(0)
    for _ in range(50):
        x = rng.integers(0, 256, 32, dtype=np.uint8); y = rng.integers(0, 256, 32, dtype=np.uint8)
        assert P.hamming(x, y) == int(np.unpackbits(x ^ y).sum())


def test_knn2_rules(P):
    """Strict '<' in ascending order: lowest index wins ties; d2 is the second smallest WITH multiplicity."""
    t, q = synth_descriptors(3000, 200, seed=4, dup=200)
    t[2000] = t[10]; q[0] = t[10]
    i1, d1, d2 = P.knn2(q, t)
    j1, e1, e2 = P.knn2(q, t, popcnt=True)
    assert np.array_equal(i1, j1) and np.array_equal(d1, e1) and np.array_equal(d2, e2)
    assert i1[0] == 10 and d1[0] == 0 and d2[0] == 0
    D = np.unpackbits(q[:, None, :] ^ t[None, :, :], axis=2).sum(2)
    assert np.array_equal(i1, D.argmin(1)) and np.array_equal(d1, D.min(1))
    assert np.array_equal(d2, np.sort(D, axis=1)[:, 1])
    # single train row: no second best
    i1, d1, d2 = P.knn2(q[:3], t[:1])
    assert list(i1) == [0, 0, 0] and list(d2) == [256, 256, 256]


def test_window_search_rules(P):
    t, q = synth_descriptors(500, 6, seed=5, dup=0)
    off = np.array([0, 0, 1, 4, 8, 8, 12], np.int32)
    idx = np.array([7, 3, 3, 9, 1, 2, 1, 2, 4, 5, 6, 4], np.int32)
    bi, bd, sd, si = P.window_search(q, off, idx, None, t)
    assert bi[0] == -1 and bd[0] == 256 and si[0] == -1          # empty list
    assert bi[1] == 7 and si[1] == -1 and sd[1] == 256           # single candidate
    assert bi[2] == 3 and sd[2] == bd[2] and si[2] == 3 or bi[2] == 9  # duplicate candidate ties with itself
    skip = np.zeros(12, np.uint8); skip[0] = 1
    bi2, bd2, _, _ = P.window_search(q, off, idx, skip, t)
    assert bi2[1] == -1 and bd2[1] == 256


def test_search_by_projection_oracle_micro_cases():
    """Hand-derived cases pin the oracle's SearchByProjection(Frame, Frame) restatement (the reference ships no test):
    two map points at the same place compete for one keypoint — the first (with observations) claims it and blocks the
    second, which falls back to the farther keypoint; without observations the first does not block and is overwritten."""
    import ctypes as C
    from oracle_lib import OraclePrims
    from orbslam2_b200 import _lib
    from orbslam2_b200.frame import Frame, forward_backward
    from orbslam2_b200.extractor import KP_DTYPE
    sf = (np.float32(1.2) ** np.arange(8, dtype=np.float32)).astype(np.float32)
    kc = np.zeros(2, KP_DTYPE); kc["x"] = (320, 323); kc["y"] = 240
    dc = np.zeros((2, 32), np.uint8); dc[1, 0] = 0x0F            # keypoint 1 is 4 bits away from the query descriptor
    cur = Frame(kc, dc, sf, 500, 500, 320, 240, 40, 640, 480)
    kl = np.zeros(2, KP_DTYPE); kl["x"] = 320; kl["y"] = 240
    last = Frame(kl, np.zeros((2, 32), np.uint8), sf, 500, 500, 320, 240, 40, 640, 480)
    last.mp_valid[:] = 1; last.mp_world[:] = (0.0, 0.0, 4.0)

    def params(th):
        P = _lib.ProjParams()
        P.fx = P.fy = 500; P.cx, P.cy, P.bf = 320, 240, 40
        P.minX, P.maxX, P.minY, P.maxY = 0, 640, 0, 480
        P.grid_w_inv, P.grid_h_inv = 64 / 640, 48 / 480
        P.nlevels = 8
        for i in range(8):
            P.scale[i] = float(sf[i])
        for i, v in enumerate(np.eye(3, dtype=np.float32).reshape(-1)):
            P.Rcw[i] = float(v)
        P.th, P.th_high, P.check_orientation = th, 100, 0
        return P

    O = OraclePrims()
    last.mp_observed[:] = 1
    n, cm, lm, lk = O.search_by_projection(params(7.0), cur, last)
    assert n == 2 and list(lm) == [0, 1] and list(cm) == [0, 1] and list(lk) == [1, 1]
    last.mp_observed[:] = 0                                         # no observations: no blocking, the later claim wins
    n, cm, lm, lk = O.search_by_projection(params(7.0), cur, last)
    assert n == 2 and list(lm) == [0, 0] and list(cm) == [1, -1]
    last.mp_observed[:] = 1
    n, cm, lm, lk = O.search_by_projection(params(2.0), cur, last)  # window 2 px: keypoint 1 (3 px away) is out of reach
    assert n == 1 and list(lm) == [0, -1] and list(cm) == [0, -1]
    assert forward_backward(cur, last, True) == (False, False)


def test_distinctive_descriptors_oracle_micro_case():
    """Three views: rows 0 and 1 identical, row 2 eight bits away. Sorted rows: [0,0,8], [0,0,8], [0,8,8]; the median is
    element int(0.5*(3-1)) = 1 -> medians 0, 0, 8 -> the first row wins. Two views: k = int(0.5) = 0 -> every median is the
    self distance 0 -> row 0."""
    from oracle_lib import OraclePrims
    d = np.zeros((5, 32), np.uint8); d[2, 0] = 0xFF; d[4, :] = 0xFF
    bi, bm = OraclePrims().distinctive_descriptors(d, np.array([0, 3, 5, 5], np.int32))
    assert list(bi) == [0, 0, -1] and list(bm[:2]) == [0, 0]


# ---- upstream image prep: restatements vs cv2 golden vectors ------------------------------------
def _prep_golden():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cv2_prep.npz"))


def colour_cube():
    r, g, b = np.meshgrid(np.arange(256), np.arange(256), np.arange(0, 256, 51), indexing="ij")
    return np.stack([r, g, b], -1).astype(np.uint8).reshape(256, -1, 3)


def test_prep_cvtcolor_matches_cv2():
    from oracle_lib import OraclePrep
    G, P = _prep_golden(), OraclePrep()
    assert np.array_equal(P.cvt_gray(G["cvt_rgb"], False), G["gray_rgb"])
    assert np.array_equal(P.cvt_gray(G["cvt_rgb"], True), G["gray_bgr"])
    assert np.array_equal(P.cvt_gray(G["cvt_rgba"], False), G["gray_rgba"])
    assert np.array_equal(P.cvt_gray(G["cvt_rgba"], True), G["gray_bgra"])
    assert np.array_equal(P.cvt_gray(colour_cube(), False), G["gray_cube_rgb"])


def test_prep_remap_matches_cv2():
    from oracle_lib import OraclePrep
    G, P = _prep_golden(), OraclePrep()
    for name in ("rect", "noisy", "wild"):
        got = P.remap(G["remap_src"], G[f"remap_{name}_x"], G[f"remap_{name}_y"])
        assert np.array_equal(got, G[f"remap_{name}"]), name


def test_prep_undistort_points_matches_cv2():
    from oracle_lib import OraclePrep
    G, P = _prep_golden(), OraclePrep()
    for name in ("tum1", "euroc4"):
        K = G[f"und_{name}_K"]
        got = P.undistort_points(G["und_pts"], K[0], K[1], K[2], K[3], G[f"und_{name}_D"])
        assert got.tobytes() == G[f"und_{name}"].tobytes(), name


def test_search_for_initialization_oracle_vs_python_replay():
    """orc_search_for_initialization against a plain-Python replay of src/ORBmatcher.cc:400-515 written independently
    (dict grid, Python ints, math.floor/ceil), on a pair built so that keypoints get stolen and late queries are refused."""
    import math
    import importlib.util, os
    from oracle_lib import oracle_search_for_initialization
    from orbslam2_b200 import _lib
    spec = importlib.util.spec_from_file_location("tgs", os.path.join(os.path.dirname(__file__), "test_gpu_search.py"))
    tgs = importlib.util.module_from_spec(spec); spec.loader.exec_module(tgs)
    F1, F2 = tgs._init_pair(11, n=400, groups=40)
    window, ratio = 100, np.float32(0.9)
    P = _lib.ProjParams()
    P.minX, P.maxX, P.minY, P.maxY = 0, 640, 0, 480
    P.grid_w_inv, P.grid_h_inv = float(F2.mfGridElementWidthInv), float(F2.mfGridElementHeightInv)
    P.nlevels, P.check_orientation = 8, 1
    prev = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
    eprev = prev.copy()
    n, m12 = oracle_search_for_initialization(P, ratio, F1, F2, prev, window)

    f32 = np.float32
    gwi, ghi = f32(P.grid_w_inv), f32(P.grid_h_inv)
    grid = {}
    for i, k in enumerate(F2.mvKeysUn):
        px = int(math.floor(float(f32(k["x"]) * gwi) + 0.5))  # round(): half away from zero, values are >= 0 here
        py = int(math.floor(float(f32(k["y"]) * ghi) + 0.5))
        if 0 <= px < 64 and 0 <= py < 48:
            grid.setdefault((px, py), []).append(i)
    bits2 = np.unpackbits(F2.mDescriptors, axis=1).astype(np.int32)
    bits1 = np.unpackbits(F1.mDescriptors, axis=1).astype(np.int32)
    INF = 2**31 - 1
    md, m21, e12 = [INF] * F2.N, [-1] * F2.N, [-1] * F1.N
    hist = [[] for _ in range(30)]
    steals = refused = nm = 0
    r = f32(window)
    for i1 in range(F1.N):
        if F1.mvKeysUn["octave"][i1] > 0:
            continue
        x, y = f32(eprev[i1, 0]), f32(eprev[i1, 1])
        x0 = max(0, math.floor(float(f32(x - r) * gwi))); x1 = min(63, math.ceil(float(f32(x + r) * gwi)))
        y0 = max(0, math.floor(float(f32(y - r) * ghi))); y1 = min(47, math.ceil(float(f32(y + r) * ghi)))
        if x0 >= 64 or x1 < 0 or y0 >= 48 or y1 < 0:
            continue
        best = best2 = INF; bi = -1
        for ix in range(x0, x1 + 1):
            for iy in range(y0, y1 + 1):
                for i2 in grid.get((ix, iy), []):
                    k = F2.mvKeysUn[i2]
                    if k["octave"] != 0 or not (abs(f32(k["x"]) - x) < r and abs(f32(k["y"]) - y) < r):
                        continue
                    d = int(np.abs(bits1[i1] - bits2[i2]).sum())
                    if md[i2] <= d:
                        refused += 1
                        continue
                    if d < best:
                        best2, best, bi = best, d, i2
                    elif d < best2:
                        best2 = d
        if best <= 50 and f32(best) < f32(f32(best2) * ratio):
            if m21[bi] >= 0:
                e12[m21[bi]] = -1; nm -= 1; steals += 1
            e12[i1] = bi; m21[bi] = i1; md[bi] = best; nm += 1
            rot = f32(F1.mvKeysUn["angle"][i1]) - f32(F2.mvKeysUn["angle"][bi])
            if rot < 0:
                rot = f32(rot + f32(360))
            b = int(math.floor(float(f32(rot * f32(1.0 / 30))) + 0.5))
            hist[0 if b == 30 else b].append(i1)
    sizes = [len(h) for h in hist]
    max1 = max2 = max3 = 0; i1_ = i2_ = i3_ = -1
    for i, s in enumerate(sizes):
        if s > max1:
            max3, max2, max1, i3_, i2_, i1_ = max2, max1, s, i2_, i1_, i
        elif s > max2:
            max3, max2, i3_, i2_ = max2, s, i2_, i
        elif s > max3:
            max3, i3_ = s, i
    if max2 < f32(0.1) * f32(max1):
        i2_ = i3_ = -1
    elif max3 < f32(0.1) * f32(max1):
        i3_ = -1
    for i, h in enumerate(hist):
        if i in (i1_, i2_, i3_):
            continue
        for q in h:
            if e12[q] >= 0:
                e12[q] = -1; nm -= 1
    assert steals >= 5 and refused >= 5, (steals, refused)
    assert n == nm and list(m12) == e12
    for q in range(F1.N):
        exp = (F2.mvKeysUn["x"][e12[q]], F2.mvKeysUn["y"][e12[q]]) if e12[q] >= 0 else tuple(eprev[q])
        assert tuple(prev[q]) == exp
