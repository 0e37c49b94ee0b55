"""GPU parity of the device-resident ORBmatcher::SearchByProjection(Frame, Frame) (grid build + GetFeaturesInArea +
Hamming scan + greedy claims + rotation filter) against the oracle's line-by-line sequential restatement.
Bit-exact bar: nmatches, the final mvpMapPoints image, every query's claim and its rotation-filter fate."""
import numpy as np
import pytest

from oracle_lib import OraclePrims
from orbslam2_b200.synth import synth_tracking_pair

pytestmark = pytest.mark.gpu


def _run(m, P, last, cur, th, bMono, taken=None):
    prm = m._proj_params(cur, last, th, bMono)
    got = m.SearchByProjection(cur, last, th, bMono, taken=taken)
    exp = P.search_by_projection(prm, cur, last, taken=taken)
    assert got[0] == exp[0], f"nmatches {got[0]} vs {exp[0]}"
    for g, e, name in zip(got[1:], exp[1:], ("cur_match", "last_match", "last_keep")):
        assert np.array_equal(g, e), f"{name}: {(g != e).sum()} differ"
    return got, prm


@pytest.mark.parametrize("seed,stereo,th,bMono,motion", [
    (0, True, 15.0, False, (0.05, -0.02, 0.12)),     # backward (stereo)
    (1, True, 7.0, False, (0.03, 0.01, -0.15)),      # forward (stereo)
    (2, True, 15.0, False, (0.02, 0.01, 0.01)),      # neither: octave window [o-1, o+1]
    (3, False, 15.0, True, (0.05, -0.02, 0.12)),     # monocular
    (4, False, 30.0, True, (0.08, 0.03, 0.10)),      # the doubled-window retry of TrackWithMotionModel
])
def test_search_by_projection_vs_oracle(seed, stereo, th, bMono, motion):
    import orbslam2_b200 as ob
    m = ob.ORBmatcher(0.9, True); P = OraclePrims()
    last, cur = synth_tracking_pair(1500, seed, stereo=stereo, motion=motion)
    got, prm = _run(m, P, last, cur, th, bMono)
    assert got[0] > 300, "synthetic pair should track"
    assert (got[2] >= 0).sum() > got[3].sum(), "the rotation filter should reject something"
    rounds, cands = m.search_stats()
    assert rounds >= 2 and cands > 1000, "conflicting claims must exercise the fixed point"
    if seed == 0:
        assert (prm.forward, prm.backward) == (0, 1)
    if seed == 1:
        assert (prm.forward, prm.backward) == (1, 0)
    if seed == 2:
        assert (prm.forward, prm.backward) == (0, 0)


def test_taken_mask_no_orientation_and_empty():
    import orbslam2_b200 as ob
    P = OraclePrims()
    last, cur = synth_tracking_pair(900, 7, width=752, height=480)
    m = ob.ORBmatcher(0.9, False)  # mbCheckOrientation = false
    taken = (np.random.default_rng(1).random(cur.N) < 0.3).astype(np.uint8)
    got, _ = _run(m, P, last, cur, 15.0, False, taken=taken)
    assert not (got[1][taken != 0] >= 0).any(), "keypoints occupied on entry are never claimed"
    assert np.array_equal(got[3], (got[2] >= 0).astype(np.uint8))
    # every map point invalid -> no match; empty current frame
    last.mp_valid[:] = 0
    got, _ = _run(m, P, last, cur, 15.0, False)
    assert got[0] == 0 and (got[1] < 0).all()


@pytest.mark.parametrize("K", [12, 70])
def test_chain_of_claims(K):
    """A worst case for the fixed point: K queries with identical descriptors all prefer keypoint 0, then 1, ... —
    query k must end on keypoint k (K rounds), exactly as the sequential loop. K = 70 goes through several of the
    doubling round groups (4 speculative rounds, then 4, 8, 16, 32, ... per host synchronisation)."""
    import orbslam2_b200 as ob
    from orbslam2_b200.frame import Frame
    from orbslam2_b200.extractor import KP_DTYPE
    sf = np.float32(1.2) ** np.arange(8, dtype=np.float32)
    kc = np.zeros(K, KP_DTYPE); kc["x"] = 320 + np.arange(K) * 0.2; kc["y"] = 240; kc["octave"] = 0
    dc = np.zeros((K, 32), np.uint8)
    for k in range(K):  # keypoint k differs from the query descriptor in k bits -> strictly increasing distance
        for b in range(k):
            dc[k, b >> 3] |= np.uint8(1 << (b & 7))
    cur = Frame(kc, dc, sf, 500, 500, 320, 240, 40, 640, 480)
    kl = np.zeros(K, KP_DTYPE); kl["x"] = 320; kl["y"] = 240
    last = Frame(kl, np.zeros((K, 32), np.uint8), sf, 500, 500, 320, 240, 40, 640, 480)
    last.mp_valid[:] = 1; last.mp_observed[:] = 1
    last.mp_world[:] = (0.0, 0.0, 5.0)
    m = ob.ORBmatcher(0.9, False); P = OraclePrims()
    got, _ = _run(m, P, last, cur, 15.0, True)
    assert list(got[2]) == list(range(K)) and got[0] == K
    assert m.search_stats()[0] >= K


def _local_points(last, cur, seed):
    """Local map points for SearchLocalPoints: the pair's map points projected with the current pose (what
    Frame::isInFrustum stores in each MapPoint), a predicted level and a viewing cosine."""
    rng = np.random.default_rng(seed)
    T = cur.mTcw.astype(np.float64)
    Xc = (T[:3, :3] @ last.mp_world.T.astype(np.float64)).T + T[:3, 3]
    u = (float(cur.fx) * Xc[:, 0] / Xc[:, 2] + float(cur.cx)).astype(np.float32)
    v = (float(cur.fy) * Xc[:, 1] / Xc[:, 2] + float(cur.cy)).astype(np.float32)
    return {"proj_x": u, "proj_y": v, "proj_xr": (u - float(cur.mbf) / Xc[:, 2]).astype(np.float32),
            "level": np.clip(last.mvKeysUn["octave"] + rng.integers(0, 2, last.N), 0, 7).astype(np.int32),
            "view_cos": rng.choice(np.array([0.9, 0.998, 0.9995], np.float32), last.N),
            "in_view": (rng.random(last.N) > 0.1) & (u > 0) & (u < 640) & (v > 0) & (v < 480),
            "observed": rng.random(last.N) > 0.1, "desc": last.mp_desc}


@pytest.mark.parametrize("seed,stereo,th,ratio", [(0, True, 1.0, 0.8), (1, True, 3.0, 0.8), (2, False, 5.0, 0.9), (3, True, 1.0, 0.6)])
def test_search_local_points_vs_oracle(seed, stereo, th, ratio):
    import orbslam2_b200 as ob
    m = ob.ORBmatcher(ratio, True); P = OraclePrims()
    last, cur = synth_tracking_pair(1500, 10 + seed, stereo=stereo, conflicts=300)
    mps = _local_points(last, cur, seed)
    taken = (np.random.default_rng(seed).random(cur.N) < 0.1).astype(np.uint8)
    n, cm, mm, prm = m.SearchLocalPoints(cur, mps, th, taken=taken)
    en, ecm, emm = P.search_local_points(prm, m.mfNNratio, cur, m._mp_arrays(mps), taken=taken)
    assert n == en and np.array_equal(cm, ecm) and np.array_equal(mm, emm)
    assert n > 200, "local points should be found"
    assert m.search_stats()[0] >= 2


@pytest.mark.parametrize("seed,th,orbdist", [(0, 10.0, 100), (1, 3.0, 64), (2, 10.0, 100)])
def test_relocalization_search_vs_oracle(seed, th, orbdist):
    """ORBmatcher::SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1468-1595) through
    the same entry point: host-predicted levels, window [l-1, l+1], no depth test, no stereo gate, any existing map
    point blocks a keypoint, accept at ORBdist."""
    import orbslam2_b200 as ob
    rng = np.random.default_rng(100 + seed)
    m = ob.ORBmatcher(0.9, True); P = OraclePrims()
    kf, cur = synth_tracking_pair(1500, 20 + seed, stereo=True, motion=(0.04, 0.02, 0.06))
    # PredictScale is the caller's job: here the keypoint's own octave jittered by -1..+1
    lvl = np.clip(kf.mvKeysUn["octave"] + rng.integers(-1, 2, kf.N), 0, 7).astype(np.int32)
    usable = ((kf.mp_valid != 0) & (rng.random(kf.N) < 0.85)).astype(np.uint8)   # bad / already found / out of range
    taken = (rng.random(cur.N) < 0.2).astype(np.uint8)                            # CurrentFrame.mvpMapPoints[i2] set
    if seed == 2:                                                                  # a pose that puts points behind the camera
        T = cur.mTcw.copy(); T[2, 3] -= np.float32(4.0); cur.SetPose(T)
    got = m.SearchByProjectionKF(cur, kf, usable, lvl, th, orbdist, taken=taken)
    prm = m._proj_params(cur, kf, th, True)
    prm.forward = prm.backward = 0
    prm.th_high, prm.skip_depth_test = orbdist, 1
    exp = P.search_by_projection(prm, cur, kf, taken=taken, flags=(usable != 0).astype(np.uint8) | np.uint8(2), octave=lvl,
                                 no_uright=True)
    assert got[0] == exp[0], f"nmatches {got[0]} vs {exp[0]}"
    for g, e, name in zip(got[1:], exp[1:], ("cur_match", "kf_match", "kf_keep")):
        assert np.array_equal(g, e), f"{name}: {(g != e).sum()} differ"
    if seed < 2:
        assert got[0] > 150
    assert not (taken[got[1] >= 0]).any(), "keypoints that already had a map point must stay untouched"
    m.close()


@pytest.mark.parametrize("seed,stereo,th,chi2", [(0, True, 3.0, True), (1, False, 3.0, True), (2, True, 4.0, False), (3, True, 9.0, True)])
def test_fuse_search_vs_oracle(seed, stereo, th, chi2):
    """Search half of ORBmatcher::Fuse (both overloads): KeyFrame grid window, octave gate, chi-square gates, best
    keypoint per map point (first wins ties) — bit-identical to the sequential restatement."""
    import orbslam2_b200 as ob
    from oracle_lib import oracle_fuse_search
    rng = np.random.default_rng(300 + seed)
    m = ob.ORBmatcher(0.6, True)
    kf, other = synth_tracking_pair(1500, 30 + seed, stereo=stereo, motion=(0.03, 0.01, 0.05))
    # map points seen from the other frame, "projected" near their true keypoints in kf (+ noise, + some far off)
    n = other.N
    src = rng.integers(0, kf.N, n)
    u = kf.mvKeysUn["x"][src] + rng.normal(0, 1.2, n).astype(np.float32)
    v = kf.mvKeysUn["y"][src] + rng.normal(0, 1.2, n).astype(np.float32)
    far = rng.random(n) < 0.15
    u[far] = rng.uniform(-30, 700, far.sum()); v[far] = rng.uniform(-30, 520, far.sum())
    if stereo:
        ur = (np.where(kf.mvuRight[src] >= 0, kf.mvuRight[src], u - 10) + rng.normal(0, 1.0, n)).astype(np.float32)
    else:
        ur = (u - rng.uniform(2, 40, n)).astype(np.float32)
    desc = kf.mDescriptors[src].copy()
    bits = np.unpackbits(desc, axis=1)
    for _ in range(8):
        bits[np.arange(n), rng.integers(0, 256, n)] ^= 1
    desc = np.packbits(bits, axis=1)
    pts = dict(u=u.astype(np.float32), v=v.astype(np.float32), ur=ur,
               level=np.clip(kf.mvKeysUn["octave"][src] + rng.integers(0, 2, n), 0, 7).astype(np.int32),
               valid=(rng.random(n) < 0.9).astype(np.uint8), desc=desc)
    bi, bd, P, inv_s2 = m.FuseSearch(kf, pts, th=th, check_chi2=chi2)
    ei, ed = oracle_fuse_search(P, inv_s2, chi2, pts, kf)
    assert np.array_equal(bi, ei) and np.array_equal(bd, ed), f"{(bi != ei).sum()} best indices differ"
    assert ((bd <= 50) & (bi >= 0)).sum() > 300, "most projected points must find their keypoint"
    assert (bi[pts["valid"] == 0] == -1).all()
    # empty sides
    e = dict(u=np.zeros(0), v=np.zeros(0), ur=np.zeros(0), level=np.zeros(0), valid=np.zeros(0), desc=np.zeros((0, 32)))
    assert len(m.FuseSearch(kf, e)[0]) == 0
    m.close()


def _projected_points(kf, n, rng, noise=1.2, flips=8, far_frac=0.15, dup_frac=0.0):
    """Map points 'projected' near keypoints of kf: position noise, descriptor bit flips, some far off; with dup_frac
    several points aim at the same keypoint (they compete for it in index order)."""
    src = rng.integers(0, kf.N, n)
    if dup_frac > 0:
        d = rng.random(n) < dup_frac
        src[d] = src[rng.integers(0, max(n // 8, 1), d.sum())]
    u = kf.mvKeysUn["x"][src] + rng.normal(0, noise, n).astype(np.float32)
    v = kf.mvKeysUn["y"][src] + rng.normal(0, noise, n).astype(np.float32)
    far = rng.random(n) < far_frac
    u[far] = rng.uniform(-30, 700, far.sum()); v[far] = rng.uniform(-30, 520, far.sum())
    bits = np.unpackbits(kf.mDescriptors[src], axis=1)
    for _ in range(flips):
        bits[np.arange(n), rng.integers(0, 256, n)] ^= 1
    return dict(u=u.astype(np.float32), v=v.astype(np.float32),
                level=np.clip(kf.mvKeysUn["octave"][src] + rng.integers(0, 2, n), 0, 7).astype(np.int32),
                valid=(rng.random(n) < 0.9).astype(np.uint8), desc=np.packbits(bits, axis=1)), src


@pytest.mark.parametrize("seed,th,dup,with_taken", [(0, 10, 0.0, True), (1, 10, 0.6, False), (2, 4, 0.3, True), (3, 25, 0.8, True)])
def test_search_kf_projection_vs_oracle(seed, th, dup, with_taken):
    """ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) (LoopClosing::ComputeSim3): KeyFrame grid
    window, octave gate, 'skip vpMatched[idx]' with the claims of earlier points — bit-identical to the sequential loop,
    also when most points compete for the same keypoints."""
    import orbslam2_b200 as ob
    from oracle_lib import oracle_search_kf_projection
    rng = np.random.default_rng(500 + seed)
    m = ob.ORBmatcher(0.75, True)
    kf, _ = synth_tracking_pair(1500, 40 + seed, stereo=bool(seed & 1), motion=(0.03, 0.01, 0.05))
    pts, _ = _projected_points(kf, 2500, rng, dup_frac=dup)
    taken = (rng.random(kf.N) < 0.2).astype(np.uint8) if with_taken else None
    n, km, pm, P = m.SearchByProjectionSim3(kf, pts, th=th, taken=taken)
    en, ekm, epm = oracle_search_kf_projection(P, pts, kf, taken)
    assert n == en and np.array_equal(km, ekm) and np.array_equal(pm, epm), f"{(pm != epm).sum()} point matches differ"
    assert n > 400
    if taken is not None:
        assert not taken[km >= 0].any() and not taken[pm[pm >= 0]].any()
    got = pm[pm >= 0]
    assert len(np.unique(got)) == len(got), "a keypoint is claimed at most once"
    # empty sides
    e = dict(u=np.zeros(0), v=np.zeros(0), level=np.zeros(0), valid=np.zeros(0), desc=np.zeros((0, 32)))
    assert m.SearchByProjectionSim3(kf, e)[0] == 0
    m.close()


@pytest.mark.parametrize("seed", [0, 1])
def test_search_by_sim3_vs_oracle(seed):
    """ORBmatcher::SearchBySim3: both window searches (octave gate, best <= TH_HIGH) and the mutual-agreement step."""
    import orbslam2_b200 as ob
    from oracle_lib import oracle_fuse_search
    rng = np.random.default_rng(600 + seed)
    m = ob.ORBmatcher(0.75, True)
    kf1, kf2 = synth_tracking_pair(1200, 60 + seed, stereo=True, motion=(0.02, 0.01, 0.03))
    # KF1's point i projected into KF2 (entry i <-> keypoint i of KF1) and the reverse; descriptors are the keypoints' own
    def side(src_kf, dst_kf):
        n = src_kf.N
        p, src = _projected_points(dst_kf, n, rng, noise=1.0, flips=0, far_frac=0.1)
        p["desc"] = np.ascontiguousarray(src_kf.mDescriptors)
        # make about half of the pairs consistent: keypoint i of src really looks like keypoint src[i] of dst
        return p, src
    p12, s12 = side(kf1, kf2)
    d1 = kf1.mDescriptors.copy(); half = rng.random(kf1.N) < 0.6
    d1[half] = kf2.mDescriptors[s12[half]]
    kf1.mDescriptors = np.ascontiguousarray(d1); p12["desc"] = kf1.mDescriptors
    p21, s21 = side(kf2, kf1)
    # the reverse projections of the consistent pairs land on their partner
    inv = np.full(kf2.N, -1); inv[s12[half]] = np.nonzero(half)[0]
    hit = inv >= 0
    p21["u"][hit] = kf1.mvKeysUn["x"][inv[hit]]; p21["v"][hit] = kf1.mvKeysUn["y"][inv[hit]]
    p21["level"][hit] = kf1.mvKeysUn["octave"][inv[hit]]
    nf, m12 = m.SearchBySim3(kf1, kf2, p12, p21, th=7.5)
    def orc(kf, pts):
        P = m._kf_grid_params(kf, 7.5)
        q = dict(pts); q["ur"] = np.zeros(len(q["u"]), np.float32)
        bi, bd = oracle_fuse_search(P, np.ones(8, np.float32), 0, q, kf)
        return np.where(bd <= 100, bi, -1)
    e1, e2 = orc(kf2, p12), orc(kf1, p21)
    exp = np.array([e1[i] if e1[i] >= 0 and e2[e1[i]] == i else -1 for i in range(kf1.N)], np.int32)
    assert np.array_equal(m12, exp) and nf == int((exp >= 0).sum())
    assert nf > 100
    m.close()


def _init_pair(seed, n=1500, groups=60, width=640, height=480):
    """Two monocular frames for SearchForInitialization: F2 = F1's keypoints moved by a small flow + noise, descriptors
    with a few flipped bits, shuffled. `groups` clusters of F1 keypoints (3-5 each, all level 0, within the window of one
    another) carry near-copies of ONE F2 descriptor with a random number of flipped bits, so later queries steal the
    keypoint from earlier, worse ones and some arrive too late (vMatchedDistance <= dist)."""
    from orbslam2_b200.frame import Frame
    from orbslam2_b200.extractor import KP_DTYPE
    rng = np.random.default_rng(900 + seed)
    sf = np.float32(1.2) ** np.arange(8, dtype=np.float32)

    def kps(x, y, o, a):
        k = np.zeros(len(x), KP_DTYPE)
        k["x"], k["y"], k["octave"], k["angle"] = x, y, o, a
        k["size"] = 31 * sf[o]; k["class_id"] = -1
        return k

    x1 = rng.uniform(20, width - 20, n).astype(np.float32); y1 = rng.uniform(20, height - 20, n).astype(np.float32)
    o1 = np.where(rng.random(n) < 0.55, 0, rng.integers(1, 8, n)).astype(np.int32)
    a1 = rng.uniform(0, 360, n).astype(np.float32)
    d1 = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    perm = rng.permutation(n)
    x2 = (x1 + 6 + rng.normal(0, 3, n)).astype(np.float32)[perm]; y2 = (y1 - 4 + rng.normal(0, 3, n)).astype(np.float32)[perm]
    o2 = np.where(rng.random(n) < 0.9, o1, rng.integers(0, 8, n)).astype(np.int32)[perm]
    rot = np.where(rng.random(n) < 0.85, 12.0, rng.uniform(0, 360, n))
    a2 = np.mod(a1 - rot + rng.normal(0, 3, n), 360).astype(np.float32)[perm]
    bits = np.unpackbits(d1, axis=1)
    for _ in range(14):
        bits[np.arange(n), rng.integers(0, 256, n)] ^= 1
    d2 = np.packbits(bits, axis=1)[perm]
    inv = np.empty(n, np.int64); inv[perm] = np.arange(n)
    # competing clusters
    heads = rng.choice(np.flatnonzero(o1 == 0), groups, replace=False)
    free = np.setdiff1d(np.arange(n), heads)
    rng.shuffle(free)
    used = 0
    for h in heads:
        k = int(rng.integers(2, 5))
        members = free[used:used + k]; used += k
        tgt = inv[h]
        o2[tgt] = 0
        for mb in members:
            x1[mb] = x1[h] + rng.normal(0, 8); y1[mb] = y1[h] + rng.normal(0, 8)
            o1[mb] = 0
            a1[mb] = a1[h]
            b = np.unpackbits(d2[tgt])
            b[rng.integers(0, 256, int(rng.integers(0, 30)))] ^= 1
            d1[mb] = np.packbits(b)
    F1 = Frame(kps(x1, y1, o1, a1), d1, sf, 520, 520, 320, 240, 0, width, height)
    F2 = Frame(kps(x2, y2, o2, a2), d2, sf, 520, 520, 320, 240, 0, width, height)
    return F1, F2


@pytest.mark.parametrize("seed,window,ratio,ori,n", [(0, 100, 0.9, True, 1500), (1, 100, 0.9, False, 1500), (2, 30, 0.7, True, 1500),
                                                     (3, 250, 1.0, True, 1500), (4, 20, 0.9, True, 30000)])
def test_search_for_initialization_vs_oracle(seed, window, ratio, ori, n):
    """ORBmatcher::SearchForInitialization (Tracking::MonocularInitialization): level-0 queries, the window over F2's
    grid, vMatchedDistance skip, ratio rule, stolen matches, rotation filter over every accept, vbPrevMatched update —
    bit-identical to the sequential loop; a second call with the updated vbPrevMatched (what the tracker does frame after
    frame) stays identical too. The 30000-keypoint case takes the kernel variant with the claim state in global memory."""
    import orbslam2_b200 as ob
    from oracle_lib import oracle_search_for_initialization
    m = ob.ORBmatcher(ratio, ori)
    F1, F2 = _init_pair(seed, n=n)
    prev = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
    eprev = prev.copy()
    for it in range(2):
        n, m12, P = m.SearchForInitialization(F1, F2, prev, window)
        en, em12 = oracle_search_for_initialization(P, np.float32(ratio), F1, F2, eprev, window)
        assert n == en, f"nmatches {n} vs {en}"
        assert np.array_equal(m12, em12), f"{(m12 != em12).sum()} matches differ"
        assert prev.tobytes() == eprev.tobytes(), "vbPrevMatched differs"
        assert n == (m12 >= 0).sum()
        assert (F1.mvKeysUn["octave"][m12 >= 0] == 0).all() and (F2.mvKeysUn["octave"][m12[m12 >= 0]] == 0).all()
        got = m12[m12 >= 0]
        assert len(np.unique(got)) == len(got), "a keypoint of F2 belongs to at most one match"
        assert n > 200
    m.close()


def test_search_for_initialization_edges():
    import orbslam2_b200 as ob
    from oracle_lib import oracle_search_for_initialization
    from orbslam2_b200.frame import Frame
    m = ob.ORBmatcher(0.9, True)
    F1, F2 = _init_pair(7, n=300, groups=10)
    E = Frame(F2.mvKeysUn[:0], F2.mDescriptors[:0], F2.mvScaleFactors, 520, 520, 320, 240, 0, 640, 480)
    prev = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
    keep = prev.copy()
    n, m12, _ = m.SearchForInitialization(F1, E, prev, 100)      # empty F2
    assert n == 0 and (m12 == -1).all() and np.array_equal(prev, keep)
    n, m12, _ = m.SearchForInitialization(E, F2, np.zeros((0, 2), np.float32), 100)  # empty F1
    assert n == 0 and len(m12) == 0
    # windows far outside the image: no candidates at all
    far = (prev + np.float32(5000)).astype(np.float32)
    efar = far.copy()
    n, m12, P = m.SearchForInitialization(F1, F2, far, 100)
    en, em12 = oracle_search_for_initialization(P, np.float32(0.9), F1, F2, efar, 100)
    assert n == en == 0 and np.array_equal(m12, em12) and np.array_equal(far, efar)
    m.close()


def test_search_by_projection_device_buffers():
    """The same entry point with ORBX_MEM_DEVICE in and out (a pipeline that keeps keypoints and descriptors on the GPU):
    identical to the host-buffer call and to the oracle; the first call of a fresh handle also exercises the
    candidate-buffer growth (k_sp_guard) on a dense pair."""
    import orbslam2_b200 as ob
    P = OraclePrims()
    for n, th in ((1200, 15.0), (6000, 30.0)):
        m = ob.ORBmatcher(0.9, True)   # fresh handle: default candidate capacity
        last, cur = synth_tracking_pair(n, 21, conflicts=n // 8)
        prm = m._proj_params(cur, last, th, False)
        got_d = m.SearchByProjection(cur, last, th, False, device=True)
        got_h = m.SearchByProjection(cur, last, th, False)
        exp = P.search_by_projection(prm, cur, last)
        for got in (got_d, got_h):
            assert got[0] == exp[0]
            for g, e, name in zip(got[1:], exp[1:], ("cur_match", "last_match", "last_keep")):
                assert np.array_equal(g, e), name
        m.close()


def test_grid_build_fused_and_kernel_sequence_agree(monkeypatch):
    """The 64 x 48 grid is built by one CTA (k_sp_grid_build) for camera-sized frames and by the count / scan / fill / sort
    kernel sequence above ORBX_GRID_FUSED_MAX keypoints: both must give the oracle's result on the same pair."""
    import orbslam2_b200 as ob
    P = OraclePrims()
    last, cur = synth_tracking_pair(2500, 12, stereo=True, motion=(0.05, -0.02, 0.12))
    fused = ob.ORBmatcher(0.9, True)
    monkeypatch.setenv("ORBX_GRID_FUSED_MAX", "0")
    seq = ob.ORBmatcher(0.9, True)
    monkeypatch.delenv("ORBX_GRID_FUSED_MAX")
    l0 = fused.launch_count(); a, _ = _run(fused, P, last, cur, 15.0, False); la = fused.launch_count() - l0
    l0 = seq.launch_count(); b, _ = _run(seq, P, last, cur, 15.0, False); lb = seq.launch_count() - l0
    assert lb - la == 3
    for x, y in zip(a[1:], b[1:]):
        assert np.array_equal(x, y)
    fused.close(); seq.close()
