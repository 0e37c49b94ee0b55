"""GPU parity: the device vocabulary-tree transform (orbv_* entry points through orbslam2_b200.Vocabulary) vs the CPU
oracle (orc_fbow_*, itself pinned against the unmodified reference fbow.cpp in tests/test_cpu_fbow.py).

Bar: bit-exact — word ids, node ids, feature lists, and the float32 weight sums / float64 scores bit for bit."""
import numpy as np
import pytest

from oracle_lib import OracleVocabulary, TUM1
from orbslam2_b200.synth import synth_image, synth_vocabulary, synth_vocabulary_features
from test_cpu_fbow import same

pytestmark = pytest.mark.gpu

VOCABS = [dict(k=10, levels=4, seed=1), dict(k=10, levels=6, seed=2, ragged=True, leaf_prob=0.15, align=32),
          dict(k=6, levels=4, seed=3, ragged=True, leaf_prob=0.1, align=16), dict(k=16, levels=2, seed=4, align=4),
          dict(k=2, levels=9, seed=5)]


@pytest.fixture(scope="module")
def ob():
    import orbslam2_b200
    return orbslam2_b200


def _features(blob, n, seed):
    rng = np.random.default_rng(seed)
    f = synth_vocabulary_features(blob, n, seed=seed)
    f[rng.random(n) < 0.05] = f[0]      # repeated descriptors: one word hit several times
    return f


@pytest.mark.parametrize("kw", VOCABS)
def test_transform_levels_vs_oracle(ob, kw):
    blob = synth_vocabulary(**kw)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    orc = OracleVocabulary(blob)
    feats = _features(blob, 1800, kw["seed"])
    for lvl in (0, 1, 2, 4, 12):
        assert same(voc.transform(feats, lvl), orc.transform(feats, lvl)), lvl
    a, b = voc.transform(feats), orc.transform(feats)
    assert np.array_equal(a[0], b[0]) and a[1].tobytes() == b[1].tobytes(), "normalised fBow"
    voc.close()


def test_ragged_batch_and_sort_sizes(ob):
    """One call over frames with 0, 1, 2, 3, 1023..1025 and 5000 features (every bitonic-sort size incl. the padding
    cases); results equal the oracle frame by frame and the single-frame call."""
    blob = synth_vocabulary(k=10, levels=5, seed=9, ragged=True, leaf_prob=0.05)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    orc = OracleVocabulary(blob)
    counts = [0, 1, 2, 3, 700, 1023, 1024, 1025, 5000, 17]
    cap = max(counts)
    desc = np.zeros((len(counts), cap, 32), np.uint8)
    for f, n in enumerate(counts):
        if n:
            desc[f, :n] = _features(blob, n, 100 + f)
    res = voc.transform_batch(desc, counts, 4)
    for f, n in enumerate(counts):
        if n == 0:
            assert len(res[f][0][0]) == 0 and res[f][1] == {}
            continue
        assert same(res[f], orc.transform(desc[f, :n], 4)), f"frame {f} ({n} features)"
    assert same(voc.transform(desc[4, :700], 4), res[4])
    with pytest.raises(RuntimeError):
        voc.transform(np.zeros((0, 32), np.uint8), 4)       # "No input data" (fbow.cpp:53)
    with pytest.raises(ob.OrbxError):
        voc.transform_batch(np.zeros((1, 20000, 32), np.uint8), [20000], 4)   # above the per-frame limit
    voc.close()


def test_getters_and_score(ob):
    blob = synth_vocabulary(k=10, levels=3, seed=11)
    voc = ob.Vocabulary(); assert not voc.isValid()
    voc.fromStream(blob)
    assert voc.isValid() and voc.getK() == 10 and voc.size() == 111 and voc.getDescSize() == 32
    assert voc.getTotalWords() == 1000
    orc = OracleVocabulary(blob)
    vecs = [voc.transform(_features(blob, n, 200 + i)) for i, n in enumerate((900, 1200, 40, 1, 1500))]
    q = vecs[0]
    got = voc.score_batch(q, vecs)
    exp = np.array([orc.score(q, v) for v in vecs])
    assert got.tobytes() == exp.tobytes()
    assert got[0] == 1.0 or abs(got[0] - 1.0) < 1e-6
    assert voc.score(vecs[1], vecs[2]) == orc.score(vecs[1], vecs[2])
    empty = (np.zeros(0, np.uint32), np.zeros(0, np.float32))
    assert voc.score(q, empty) == orc.score(q, empty) == 0.0
    voc.close(); assert not voc.isValid()


def test_device_descriptors_from_extractor(ob):
    """Frame::ComputeFboW on the extractor's device output: the descriptors go from orbx_extract_batch to
    orbv_transform_batch without leaving the GPU and give the result of the host path."""
    import torch
    blob = synth_vocabulary(k=10, levels=4, seed=12)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    orc = OracleVocabulary(blob)
    ex = ob.ORBextractor(**TUM1)
    imgs = torch.from_numpy(np.stack([synth_image(640, 480, s) for s in range(3)])).cuda()
    kps, desc, n = ex.extract_batch(imgs)
    res = voc.transform_batch(desc, n, 4)
    dh = desc.cpu().numpy()
    for f in range(3):
        assert same(res[f], orc.transform(dh[f, :n[f]], 4))
    assert same(voc.transform(dh[1, :n[1]], 4), res[1])
    assert voc.launch_count() > 0
    ex.close(); voc.close()


# ---- ORBmatcher::SearchByFboW on the device ---------------------------------------------------------------------------
def _bow_frames(ob, voc, blob, n_a, n_b, seed, level=4, shared=0.7):
    """two frames that see partly the same words: B's descriptors are A's with a few bits flipped (shuffled), plus
    unrelated ones; keypoint angles consistent up to a common rotation + noise (so the rotation filter has a peak)."""
    rng = np.random.default_rng(seed)
    da = _features(blob, n_a, seed)
    ns = int(min(n_a, n_b) * shared)
    pick = rng.permutation(n_a)[:ns]
    db = np.concatenate([da[pick], _features(blob, n_b - ns, seed + 1)])
    bits = np.unpackbits(db[:ns], axis=1)
    for _ in range(6):
        bits[np.arange(ns), rng.integers(0, 256, ns)] ^= 1
    db[:ns] = np.packbits(bits, axis=1)
    perm = rng.permutation(n_b)
    db = np.ascontiguousarray(db[perm])
    ang_a = (rng.random(n_a) * 360).astype(np.float32)
    ang_b = (rng.random(n_b) * 360).astype(np.float32)
    src = np.full(n_b, -1); src[:ns] = pick; src = src[perm]
    has = src >= 0
    ang_b[has] = np.mod(ang_a[src[has]] - np.float32(25.0) + rng.normal(0, 3, has.sum()).astype(np.float32), 360).astype(np.float32)
    wrong = rng.random(n_b) < 0.15
    ang_b[wrong] = (rng.random(wrong.sum()) * 360).astype(np.float32)
    return da, ang_a, None, db, ang_b, None


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("strict", [False, True])
@pytest.mark.parametrize("level", [4, 2])
def test_search_by_bow_vs_oracle(ob, strict, level, fused, monkeypatch):
    """fused: the one-CTA kernel (k_bow_fused, feature vectors up to 4096 entries); not fused: the multi-launch rounds
    that larger frames take (forced here with ORBX_BOW_FUSED_MAX=0, read when the matcher handle is created)."""
    from oracle_lib import oracle_search_by_bow
    if not fused:
        monkeypatch.setenv("ORBX_BOW_FUSED_MAX", "0")
    blob = synth_vocabulary(k=10, levels=5, seed=31)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    rng = np.random.default_rng(7)
    total = 0
    for seed, (na, nb) in enumerate([(1000, 1100), (1500, 700), (300, 2000), (5, 5)]):
        da, ang_a, _, db, ang_b, _ = _bow_frames(ob, voc, blob, na, nb, 40 + seed, level)
        fa, fb = voc.transform(da, level)[1], voc.transform(db, level)[1]
        elig_a = (rng.random(na) < 0.8).astype(np.uint8)
        elig_b = (rng.random(nb) < 0.85).astype(np.uint8) if strict else None
        A = ob.ORBmatcher.bow_side(da, ang_a, fa, elig_a)
        B = ob.ORBmatcher.bow_side(db, ang_b, fb, elig_b)
        for ratio, ori in ((0.7, True), (0.9, True), (0.75, False)):
            m = ob.ORBmatcher(ratio, ori)
            l0 = m.launch_count()
            n, match = m.SearchByFboW(A, B, strict=strict)
            if na > 5:
                assert (m.launch_count() - l0 == 1) == fused
            en, ematch = oracle_search_by_bow(A, B, ratio, 50, strict, ori)
            assert n == en and np.array_equal(match, ematch), (seed, ratio, ori)
            if ratio == 0.9:  # the same call with device buffers in and out
                nd, md = m.SearchByFboW(A, B, strict=strict, device=True)
                assert nd == en and np.array_equal(md, ematch), ("device", seed)
            total += n
            m.close()
    assert total > 500, "the synthetic frames must actually match"
    voc.close()


def test_search_by_bow_greedy_chain_and_empty(ob):
    """Adversarial order dependence: every A feature's best candidate is the same B feature, so each query's result
    depends on all earlier ones (the fixed point needs many rounds); and the empty / disjoint cases."""
    from oracle_lib import oracle_search_by_bow
    rng = np.random.default_rng(3)
    n = 40
    base = rng.integers(0, 256, 32, dtype=np.uint8)
    # B: features at Hamming distance 0, 2, 4, ... from base (distinct); A: all equal to base, so query r takes B[r]
    db = np.repeat(base[None], n, 0)
    bits = np.unpackbits(db, axis=1)
    for r in range(n):
        bits[r, :2 * r] ^= 1
    db = np.packbits(bits, axis=1)
    da = np.repeat(base[None], n, 0)
    fv_a = {7: np.arange(n, dtype=np.uint32)}
    fv_b = {7: rng.permutation(n).astype(np.uint32)}
    A = ob.ORBmatcher.bow_side(da, np.zeros(n, np.float32), fv_a, np.ones(n, np.uint8))
    B = ob.ORBmatcher.bow_side(db, np.zeros(n, np.float32), fv_b, None)
    m = ob.ORBmatcher(0.99, False)
    got = m.SearchByFboW(A, B)
    exp = oracle_search_by_bow(A, B, 0.99, 50, False, False)
    assert got[0] == exp[0] and np.array_equal(got[1], exp[1]) and got[0] >= 10
    rounds, _ = m.search_stats()
    assert rounds > 4
    # disjoint node sets and empty sides
    B2 = ob.ORBmatcher.bow_side(db, np.zeros(n, np.float32), {9: np.arange(n, dtype=np.uint32)}, None)
    assert m.SearchByFboW(A, B2)[0] == 0
    E = ob.ORBmatcher.bow_side(np.zeros((0, 32), np.uint8), np.zeros(0, np.float32), {}, None)
    assert m.SearchByFboW(A, E)[0] == 0 and m.SearchByFboW(E, B)[0] == 0
    m.close()


def _two_view_geometry(n_a, n_b, src_of_b, seed):
    """keypoints of two views of random 3-D points (B's feature j sees the point of A's feature src_of_b[j] when >= 0),
    with the fundamental matrix F12 and KF1's epipole in KF2, all float32 as the reference holds them."""
    rng = np.random.default_rng(seed)
    fx = fy = np.float32(520.0); cx, cy = np.float32(320.0), np.float32(240.0)
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float64)
    ang = 0.05
    R21 = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    t21 = np.array([0.25, 0.02, 0.05])
    X = np.stack([rng.uniform(-2, 2, n_a), rng.uniform(-1.5, 1.5, n_a), rng.uniform(3, 9, n_a)], 1)
    pa = (K @ X.T).T; pa = pa[:, :2] / pa[:, 2:]
    pt_b = np.stack([rng.uniform(0, 640, n_b), rng.uniform(0, 480, n_b)], 1)
    has = src_of_b >= 0
    Xb = (R21 @ X[src_of_b[has]].T).T + t21
    pb = (K @ Xb.T).T; pb = pb[:, :2] / pb[:, 2:]
    pt_b[has] = pb + rng.normal(0, 0.6, pb.shape)
    tx = np.array([[0, -t21[2], t21[1]], [t21[2], 0, -t21[0]], [-t21[1], t21[0], 0]])
    E = tx @ R21
    F21 = np.linalg.inv(K).T @ E @ np.linalg.inv(K)      # x2' F21 x1 = 0
    F12 = F21.T                                          # line in image 2: l = x1' F12
    C2 = t21                                             # KF1's centre in KF2's frame
    ex, ey = fx * np.float32(C2[0]) / np.float32(C2[2]) + cx, fy * np.float32(C2[1]) / np.float32(C2[2]) + cy
    sf = (np.float32(1.2) ** np.arange(8)).astype(np.float32)
    return dict(pt_a=pa.astype(np.float32), pt_b=pt_b.astype(np.float32), octave_b=rng.integers(0, 8, n_b).astype(np.int32),
                stereo_a=(rng.random(n_a) < 0.4).astype(np.uint8), stereo_b=(rng.random(n_b) < 0.4).astype(np.uint8),
                F12=F12.astype(np.float32), ex=ex, ey=ey, scale_factors=sf, level_sigma2=(sf * sf).astype(np.float32))


@pytest.mark.parametrize("only_stereo", [False, True])
def test_search_for_triangulation_vs_oracle(ob, only_stereo):
    from oracle_lib import oracle_search_for_triangulation
    blob = synth_vocabulary(k=10, levels=5, seed=33)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    total = 0
    for seed, (na, nb) in enumerate([(1200, 1300), (800, 2000), (40, 30)]):
        rng = np.random.default_rng(60 + seed)
        da = _features(blob, na, 70 + seed)
        ns = int(min(na, nb) * 0.7)
        pick = rng.permutation(na)[:ns]
        db = np.concatenate([da[pick], _features(blob, nb - ns, 80 + seed)])
        bits = np.unpackbits(db[:ns], axis=1)
        for _ in range(5):
            bits[np.arange(ns), rng.integers(0, 256, ns)] ^= 1
        db[:ns] = np.packbits(bits, axis=1)
        perm = rng.permutation(nb)
        db = np.ascontiguousarray(db[perm])
        src = np.full(nb, -1); src[:ns] = pick; src = src[perm]
        geo = _two_view_geometry(na, nb, src, seed)
        ang_a = (rng.random(na) * 360).astype(np.float32)
        ang_b = (rng.random(nb) * 360).astype(np.float32)
        ang_b[src >= 0] = np.mod(ang_a[src[src >= 0]] + np.float32(12.0) + rng.normal(0, 3, (src >= 0).sum()), 360).astype(np.float32)
        fa, fb = voc.transform(da, 4)[1], voc.transform(db, 4)[1]
        ea = (rng.random(na) < 0.7).astype(np.uint8)      # features without a map point
        eb = (rng.random(nb) < 0.7).astype(np.uint8)
        if only_stereo:
            ea &= geo["stereo_a"]; eb &= geo["stereo_b"]
        A = ob.ORBmatcher.bow_side(da, ang_a, fa, ea)
        B = ob.ORBmatcher.bow_side(db, ang_b, fb, eb)
        for ori in (True, False):
            m = ob.ORBmatcher(0.6, ori)
            n, match = m.SearchForTriangulation(A, B, geo)
            en, ematch = oracle_search_for_triangulation(A, B, geo, 50, ori)
            assert n == en and np.array_equal(match, ematch), (seed, ori)
            total += n
            m.close()
    assert total > (100 if only_stereo else 400)
    voc.close()


@pytest.mark.parametrize("strict", [False, True])
def test_search_by_bow_batch_vs_oracle(ob, strict):
    """orbm_search_by_bow_batch: K candidate keyframes against one frame (Relocalization) / keyframe (LoopClosing) in
    one launch sequence — every pair bit-identical to the sequential restatement, incl. ragged sizes, an adversarial
    chain pair that needs many rounds while the others converged long ago, a disjoint pair and empty sides."""
    from oracle_lib import oracle_search_by_bow
    blob = synth_vocabulary(k=10, levels=5, seed=33)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    rng = np.random.default_rng(17)
    nb = 1200
    pairs = []
    B = None
    for k, na in enumerate([900, 1500, 300, 1100, 40, 700]):
        da, ang_a, _, db, ang_b, _ = _bow_frames(ob, voc, blob, na, nb, 70 + k, 4)
        fa = voc.transform(da, 4)[1]
        elig_a = (rng.random(na) < 0.8).astype(np.uint8)
        A = ob.ORBmatcher.bow_side(da, ang_a, fa, elig_a)
        if strict or B is None:  # the relocalisation shape shares ONE frame among the pairs
            fb = voc.transform(db, 4)[1]
            elig_b = (rng.random(nb) < 0.85).astype(np.uint8) if strict else None
            Bk = ob.ORBmatcher.bow_side(db, ang_b, fb, elig_b)
            if B is None:
                B = Bk
        pairs.append((A, Bk if strict else B))
    # adversarial chain, disjoint nodes, empty sides
    n = 40
    base = rng.integers(0, 256, 32, dtype=np.uint8)
    bits = np.unpackbits(np.repeat(base[None], n, 0), axis=1)
    for r in range(n):
        bits[r, :2 * r] ^= 1
    dbc = np.packbits(bits, axis=1)
    dac = np.repeat(base[None], n, 0)
    Ac = ob.ORBmatcher.bow_side(dac, np.zeros(n, np.float32), {7: np.arange(n, dtype=np.uint32)}, np.ones(n, np.uint8))
    Bc = ob.ORBmatcher.bow_side(dbc, np.zeros(n, np.float32), {7: rng.permutation(n).astype(np.uint32)},
                                np.ones(n, np.uint8) if strict else None)
    Bd = ob.ORBmatcher.bow_side(dbc, np.zeros(n, np.float32), {9: np.arange(n, dtype=np.uint32)}, None)
    E = ob.ORBmatcher.bow_side(np.zeros((0, 32), np.uint8), np.zeros(0, np.float32), {}, None)
    pairs += [(Ac, Bc), (Ac, Bd), (Ac, E), (E, Bc)]
    for ratio, ori in ((0.99, False), (0.75, True)):
        m = ob.ORBmatcher(ratio, ori)
        got = m.SearchByFboWBatch(pairs, strict=strict)
        assert len(got) == len(pairs)
        total = 0
        for k, (A, Bk) in enumerate(pairs):
            en, ematch = oracle_search_by_bow(A, Bk, ratio, 50, strict, ori)
            assert got[k][0] == en and np.array_equal(got[k][1], ematch), (k, ratio, ori)
            one = m.SearchByFboW(A, Bk, strict=strict)
            assert one[0] == en and np.array_equal(one[1], ematch)
            total += en
        assert total > 500
        if not ori:
            assert got[6][0] >= 10
        assert m.SearchByFboWBatch([], strict=strict) == []
        m.close()
    voc.close()


def test_search_for_triangulation_batch_vs_oracle(ob):
    """orbm_search_for_triangulation_batch: the new keyframe against K neighbours (LocalMapping::CreateNewMapPoints) in one
    launch sequence; the keyframe's own arrays are shared by the pairs. Every pair equals the sequential restatement and
    the single-pair call."""
    from oracle_lib import oracle_search_for_triangulation
    blob = synth_vocabulary(k=10, levels=5, seed=33)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    na = 1200
    rng = np.random.default_rng(91)
    da = _features(blob, na, 170)
    ang_a = (rng.random(na) * 360).astype(np.float32)
    fa = voc.transform(da, 4)[1]
    ea = (rng.random(na) < 0.7).astype(np.uint8)
    A = ob.ORBmatcher.bow_side(da, ang_a, fa, ea)
    items, geo0 = [], None
    for k, nb in enumerate([1300, 900, 2000, 30, 1100]):
        ns = int(min(na, nb) * 0.7)
        pick = rng.permutation(na)[:ns]
        db = np.concatenate([da[pick], _features(blob, nb - ns, 180 + k)])
        bits = np.unpackbits(db[:ns], axis=1)
        for _ in range(5):
            bits[np.arange(ns), rng.integers(0, 256, ns)] ^= 1
        db[:ns] = np.packbits(bits, axis=1)
        perm = rng.permutation(nb)
        db = np.ascontiguousarray(db[perm])
        src = np.full(nb, -1); src[:ns] = pick; src = src[perm]
        geo = _two_view_geometry(na, nb, src, 5)
        if geo0 is None:
            geo0 = geo
        geo["pt_a"], geo["stereo_a"] = geo0["pt_a"], geo0["stereo_a"]
        ang_b = (rng.random(nb) * 360).astype(np.float32)
        ang_b[src >= 0] = np.mod(ang_a[src[src >= 0]] + np.float32(12.0) + rng.normal(0, 3, (src >= 0).sum()), 360).astype(np.float32)
        fb = voc.transform(db, 4)[1]
        eb = (rng.random(nb) < 0.7).astype(np.uint8)
        items.append((A, ob.ORBmatcher.bow_side(db, ang_b, fb, eb), geo))
    for ori in (True, False):
        m = ob.ORBmatcher(0.6, ori)
        got = m.SearchForTriangulationBatch(items)
        total = 0
        for k, (a, b, geo) in enumerate(items):
            en, ematch = oracle_search_for_triangulation(a, b, geo, 50, ori)
            assert got[k][0] == en and np.array_equal(got[k][1], ematch), (k, ori)
            one = m.SearchForTriangulation(a, b, geo)
            assert one[0] == en and np.array_equal(one[1], ematch)
            total += en
        assert total > 600
        assert m.SearchForTriangulationBatch([]) == []
        m.close()
    voc.close()
