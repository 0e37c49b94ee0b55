"""GPU test of the C++ drop-in wrappers (include/dropin/ORBextractor.h, ORBmatcher.h): a C++ program written like the
reference's callers (Frame::ExtractORB, stereo Frame ctor) is compiled with g++ against the cv shim, linked to
liborbx.so, run on the device, and its outputs are compared with the Python host mirror and the CPU oracle."""
import os
import subprocess

import numpy as np
import pytest

from oracle_lib import Oracle, OraclePrims, OracleVocabulary, canonical, KP_DTYPE, TUM1
from orbslam2_b200.synth import synth_stereo_pair, synth_vocabulary

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_dropin_end_to_end(tmp_path):
    from orbslam2_b200 import build
    lib = build.build()
    exe = str(tmp_path / "dropin_run")
    subprocess.check_call(["g++", "-std=c++14", "-O1", "-I", os.path.join(ROOT, "include"), "-I",
                           os.path.join(ROOT, "oracle", "shim"), os.path.join(ROOT, "tests", "cpp", "wrapper_compile_check.cpp"),
                           "-L", os.path.dirname(lib), "-lorbx", "-Wl,-rpath," + os.path.dirname(lib), "-o", exe])
    w, h = 640, 480
    left, right = synth_stereo_pair(w, h, 5)
    left.tofile(tmp_path / "l.raw"); right.tofile(tmp_path / "r.raw")
    out, out2, vfile = tmp_path / "out.bin", tmp_path / "out2.bin", tmp_path / "synthetic.fbow"
    out3, out4 = tmp_path / "out3.bin", tmp_path / "out4.bin"
    blob = synth_vocabulary(k=10, levels=4, seed=21)
    vfile.write_bytes(blob)
    r = subprocess.run([exe, str(w), str(h), str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(out), str(vfile),
                        str(out2), str(out3), str(out4)], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = open(out, "rb").read()
    nL, nR, accepted, d01, levels, p3w = np.frombuffer(raw[:24], np.int32)
    off = 24
    kps = np.frombuffer(raw[off:off + 28 * nL], KP_DTYPE); off += 28 * nL
    desc = np.frombuffer(raw[off:off + 32 * nL], np.uint8).reshape(nL, 32); off += 32 * nL
    uR = np.frombuffer(raw[off:off + 4 * nL], np.float32); off += 4 * nL
    depth = np.frombuffer(raw[off:off + 4 * nL], np.float32); off += 4 * nL
    i1 = np.frombuffer(raw[off:off + 4 * nL], np.int32); off += 4 * nL

    orc = Oracle(**TUM1); P = OraclePrims()
    ko, do = canonical(*orc.extract(left, 4000))
    assert levels == 8 and nL == len(ko)
    assert kps.tobytes() == ko.tobytes() or (np.abs(kps["x"] - ko["x"]).max() <= 1e-4 and np.abs(kps["angle"] - ko["angle"]).max() <= 1e-2)
    assert (desc == do).all(axis=1).mean() >= 0.9999
    # mvImagePyramid[3]: ROI inside a padded reflect-101 buffer (the row above the ROI is the reflection of row 1)
    lw, lh = orc.level_size(3)
    assert p3w == lw
    pyr = np.frombuffer(raw[off:off + lw * (lh + 1)], np.uint8).reshape(lh + 1, lw)
    lvl = orc.level(3)
    assert np.array_equal(pyr[1:], lvl) and np.array_equal(pyr[0], lvl[1])
    # stereo + matcher against the oracle on the same inputs
    orcR = Oracle(**TUM1)
    kr, dr = canonical(*orcR.extract(right, 4000))
    assert nR == len(kr)
    pyrL = [orc.level(l) for l in range(8)]; pyrR = [orcR.level(l) for l in range(8)]
    t, _, _ = orc.tables()
    eu, ez, _, _ = P.stereo_match(ko, do, kr, dr, pyrL, pyrR, t[0], t[1], 40.0, 0.0, 400.0)
    assert np.array_equal(uR, eu) and np.array_equal(depth, ez)
    o1, e1, e2 = P.knn2(do, dr)
    assert np.array_equal(i1, o1)
    exp_acc = int(((e1 <= 50) & (e1.astype(np.float32) < np.float32(0.7) * e2.astype(np.float32))).sum())
    assert accepted == exp_acc
    assert d01 == P.hamming(do[0], do[1])
    # fbow drop-in: Vocabulary::transform(features, 4, fBow, fBow2), transform(features) and fBow::score
    raw2 = open(out2, "rb").read()
    nw, nn, k, words = np.frombuffer(raw2[:16], np.int32)
    assert k == 10 and words == 10000
    voc = OracleVocabulary(blob)
    (ewid, eww), efv = voc.transform(desc, 4)
    pairs = np.frombuffer(raw2[16:16 + 8 * nw], np.uint32).reshape(nw, 2)
    assert np.array_equal(pairs[:, 0], ewid) and pairs[:, 1].tobytes() == eww.tobytes()
    at = 16 + 8 * nw
    assert nn == len(efv)
    for key in sorted(efv):
        nid, c = np.frombuffer(raw2[at:at + 8], np.uint32)
        lst = np.frombuffer(raw2[at + 8:at + 8 + 4 * c], np.uint32)
        assert nid == key and np.array_equal(lst, efv[key])
        at += 8 + 4 * c
    sc = np.frombuffer(raw2[at:at + 16], np.float64)
    a, bR = voc.transform(desc), voc.transform(dr)  # (right descriptors: the oracle's own extraction, identical above)
    assert sc[0] == voc.score(a, bR) and sc[1] == voc.score(a, a)

    # SearchForInitialization drop-in (Tracking::MonocularInitialization: ORBmatcher(0.9, true), window 100) on the two
    # extracted frames, against the oracle's sequential restatement on the very same keypoints and descriptors
    from oracle_lib import oracle_search_for_initialization
    from orbslam2_b200 import _lib
    import orbslam2_b200 as ob
    raw3 = open(out3, "rb").read()
    n1, n2, nm = np.frombuffer(raw3[:12], np.int32)
    at = 12
    m12 = np.frombuffer(raw3[at:at + 4 * n1], np.int32); at += 4 * n1
    prev = np.frombuffer(raw3[at:at + 8 * n1], np.float32).reshape(n1, 2); at += 8 * n1
    kR = np.frombuffer(raw3[at:at + 28 * n2], KP_DTYPE); at += 28 * n2
    dR = np.frombuffer(raw3[at:at + 32 * n2], np.uint8).reshape(n2, 32)
    assert n1 == nL and n2 == nR
    sf = np.asarray(t[0], np.float32)
    F1 = ob.Frame(kps, desc, sf, 520, 520, 320, 240, 0, w, h)
    F2 = ob.Frame(kR, dR, sf, 520, 520, 320, 240, 0, w, h)
    Pp = _lib.ProjParams()
    Pp.minX, Pp.maxX, Pp.minY, Pp.maxY = 0, w, 0, h
    Pp.grid_w_inv, Pp.grid_h_inv = float(F2.mfGridElementWidthInv), float(F2.mfGridElementHeightInv)
    Pp.nlevels, Pp.check_orientation = 8, 1
    eprev = np.ascontiguousarray(np.stack([kps["x"], kps["y"]], 1), np.float32)
    en, em12 = oracle_search_for_initialization(Pp, np.float32(0.9), F1, F2, eprev, 100)
    assert nm == en and np.array_equal(m12, em12) and prev.tobytes() == eprev.tobytes()
    assert nm > 10, "the stereo pair must give initialisation matches"

    # SearchByFboW(KeyFrame, Frame) drop-in (TrackReferenceKeyFrame: ORBmatcher(0.7, true)) on the same two frames and the
    # feature vectors the fbow drop-in produced, against the oracle's sequential restatement
    from oracle_lib import oracle_search_by_bow
    raw4 = np.fromfile(out4, np.int32)
    _, efvR = voc.transform(dR, 4)
    elig = (np.arange(nL) % 5 != 4).astype(np.uint8)
    A = ob.ORBmatcher.bow_side(desc, np.ascontiguousarray(kps["angle"]), efv, elig)
    B = ob.ORBmatcher.bow_side(dR, np.ascontiguousarray(kR["angle"]), efvR, None)
    en, ematch = oracle_search_by_bow(A, B, 0.7, 50, False, True)
    expF = np.full(n2, -1, np.int32)
    expF[ematch[ematch >= 0]] = np.flatnonzero(ematch >= 0)
    assert raw4[0] == en and np.array_equal(raw4[1:], expF), f"{(raw4[1:] != expF).sum()} frame keypoints differ"
    assert en > 5

@pytest.mark.parametrize("stereo,bMono,th", [(True, False, 15.0), (False, True, 7.0)])
def test_cpp_tracking_search_dropin(tmp_path, stereo, bMono, th):
    """The C++ template ORBmatcher::SearchByProjection(Frame&, const Frame&, th, bMono) (what Tracking.cc calls) run for
    real on a dumped frame pair: its host half (pose algebra, flags, flattening, write-back of mvpMapPoints) plus the
    device search must give the oracle's nmatches and final mvpMapPoints."""
    from orbslam2_b200 import build
    from orbslam2_b200.synth import synth_tracking_pair
    import orbslam2_b200 as ob
    lib = build.build()
    exe = str(tmp_path / "dropin_run")
    subprocess.check_call(["g++", "-std=c++14", "-O1", "-I", os.path.join(ROOT, "include"), "-I",
                           os.path.join(ROOT, "oracle", "shim"), os.path.join(ROOT, "tests", "cpp", "wrapper_compile_check.cpp"),
                           "-L", os.path.dirname(lib), "-lorbx", "-Wl,-rpath," + os.path.dirname(lib), "-o", exe])
    last, cur = synth_tracking_pair(1500, 33, stereo=stereo, motion=(0.05, -0.02, 0.12), conflicts=200)
    with open(tmp_path / "pair.bin", "wb") as f:
        f.write(np.array([last.N, cur.N, int(stereo), int(bMono)], np.int32).tobytes())
        f.write(np.array([th, cur.fx, cur.fy, cur.cx, cur.cy, cur.mbf, cur.mnMaxX, cur.mnMaxY], np.float32).tobytes())
        f.write(np.ascontiguousarray(cur.mTcw, np.float32).tobytes()); f.write(np.ascontiguousarray(last.mTcw, np.float32).tobytes())
        f.write(np.ascontiguousarray(cur.mvScaleFactors[:8], np.float32).tobytes())
        f.write(np.ascontiguousarray(last.mp_world, np.float32).tobytes())
        for a in (last.mp_valid, last.mvbOutlier, last.mp_observed):
            f.write(np.ascontiguousarray(a, np.uint8).tobytes())
        f.write(last.mvKeysUn.tobytes()); f.write(np.ascontiguousarray(last.mp_desc, np.uint8).tobytes())
        f.write(cur.mvKeysUn.tobytes())
        if stereo:
            f.write(np.ascontiguousarray(cur.mvuRight, np.float32).tobytes())
        f.write(cur.mDescriptors.tobytes())
    r = subprocess.run([exe, "track", str(tmp_path / "pair.bin"), str(tmp_path / "track.bin")], capture_output=True, text=True,
                       timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = np.fromfile(tmp_path / "track.bin", np.int32)
    m = ob.ORBmatcher(0.9, True)
    prm = m._proj_params(cur, last, th, bMono)
    exp = OraclePrims().search_by_projection(prm, cur, last)
    m.close()
    assert raw[0] == exp[0] and np.array_equal(raw[1:], exp[1]), f"{(raw[1:] != exp[1]).sum()} keypoints differ"
    assert raw[0] > 300


def _build_exe(tmp_path):
    from orbslam2_b200 import build
    lib = build.build()
    exe = str(tmp_path / "dropin_run")
    subprocess.check_call(["g++", "-std=c++14", "-O1", "-I", os.path.join(ROOT, "include"), "-I",
                           os.path.join(ROOT, "oracle", "shim"), os.path.join(ROOT, "tests", "cpp", "wrapper_compile_check.cpp"),
                           "-L", os.path.dirname(lib), "-lorbx", "-Wl,-rpath," + os.path.dirname(lib), "-o", exe])
    return exe


def test_cpp_sim3_projection_dropin(tmp_path):
    """The C++ template ORBmatcher::SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) (LoopClosing::ComputeSim3)
    run for real: Scw decomposition, projection and the MapPoint gates on the host (restated here in numpy with the same
    float / double steps), window search with the vpMatched claims on the device — against the oracle's sequential loop."""
    from orbslam2_b200.synth import synth_tracking_pair
    from oracle_lib import oracle_search_kf_projection
    from orbslam2_b200 import _lib
    exe = _build_exe(tmp_path)
    rng = np.random.default_rng(77)
    kf, _ = synth_tracking_pair(1500, 44, stereo=False, motion=(0.03, 0.01, 0.05))
    w, h, th = 640, 480, 10
    f32, f64 = np.float32, np.float64
    # Sim3 = scale * [R | t] with a small rotation; world points chosen so that they project near keypoints of kf
    s_true, ang = f32(1.7), 0.04
    R = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    t = np.array([0.1, -0.05, 0.2])
    Scw = np.eye(4, dtype=f32); Scw[:3, :3] = (s_true * R).astype(f32); Scw[:3, 3] = (s_true * t).astype(f32)
    n = 2500
    src = rng.integers(0, kf.N, n)
    z = rng.uniform(2, 10, n)
    pc = np.stack([(kf.mvKeysUn["x"][src] + rng.normal(0, 1.2, n) - float(kf.cx)) / float(kf.fx) * z,
                   (kf.mvKeysUn["y"][src] + rng.normal(0, 1.2, n) - float(kf.cy)) / float(kf.fy) * z, z], 1)
    behind = rng.random(n) < 0.05
    pc[behind, 2] *= -1
    world = ((pc - t) @ R).astype(f32)                       # Rcw^T (pc - tcw)
    bits = np.unpackbits(kf.mDescriptors[src], axis=1)
    for _ in range(8):
        bits[np.arange(n), rng.integers(0, 256, n)] ^= 1
    desc = np.packbits(bits, axis=1)
    level = np.clip(kf.mvKeysUn["octave"][src] + rng.integers(0, 2, n), 0, 7).astype(np.int32)
    bad = (rng.random(n) < 0.05).astype(np.uint8)
    dmin = rng.uniform(0.5, 3.0, n).astype(f32); dmax = rng.uniform(6.0, 14.0, n).astype(f32)
    # restatement of project_sim3 (include/dropin/ORBmatcher.h) = src/ORBmatcher.cc:293-298, 313-351
    sR = Scw[:3, :3]
    scw = f32(np.sqrt(sum(f64(sR[0, c]) * f64(sR[0, c]) for c in range(3))))
    inv = f32(1.0 / f64(scw))
    Rcw = (sR * inv).astype(f32); tcw = (Scw[:3, 3] * inv).astype(f32)
    Ow = np.array([-f32(sum(f64(Rcw[c, r]) * f64(tcw[c]) for c in range(3))) for r in range(3)], f32)
    normal = np.zeros((n, 3), f32)
    u = np.zeros(n, f32); v = np.zeros(n, f32); valid = np.zeros(n, np.uint8)
    for i in range(n):
        wv = world[i]
        po = (wv - Ow).astype(f32)
        nv = po / np.linalg.norm(po)
        if rng.random() < 0.1:
            nv = -nv                                            # viewing-angle gate fails
        normal[i] = nv.astype(f32)
    for i in range(n):
        if bad[i]:
            continue
        wv = world[i]
        p = [f32(f32(sum(f64(Rcw[r, c]) * f64(wv[c]) for c in range(3))) + tcw[r]) for r in range(3)]
        if p[2] < 0:
            continue
        invz = f32(1.0 / f64(p[2]))
        uu = f32(f32(kf.fx * f32(p[0] * invz)) + kf.cx); vv = f32(f32(kf.fy * f32(p[1] * invz)) + kf.cy)
        if not (0 <= uu < w and 0 <= vv < h):
            continue
        po = [f32(wv[k] - Ow[k]) for k in range(3)]
        dist = f32(np.sqrt(sum(f64(x) * f64(x) for x in po)))
        if dist < dmin[i] or dist > dmax[i]:
            continue
        if sum(f64(po[k]) * f64(normal[i, k]) for k in range(3)) < 0.5 * f64(dist):
            continue
        u[i], v[i], valid[i] = uu, vv, 1
    taken = (rng.random(kf.N) < 0.15).astype(np.uint8)
    with open(tmp_path / "sim3.bin", "wb") as f:
        f.write(np.array([n, kf.N, th], np.int32).tobytes())
        f.write(np.array([kf.fx, kf.fy, kf.cx, kf.cy, w, h], f32).tobytes())
        f.write(Scw.tobytes()); f.write(np.ascontiguousarray(kf.mvScaleFactors[:8], f32).tobytes())
        for a in (world, normal, dmin, dmax, level, bad, desc):
            f.write(np.ascontiguousarray(a).tobytes())
        f.write(kf.mvKeysUn.tobytes()); f.write(kf.mDescriptors.tobytes()); f.write(taken.tobytes())
    r = subprocess.run([exe, "sim3", str(tmp_path / "sim3.bin"), str(tmp_path / "sim3_out.bin")], capture_output=True, text=True,
                       timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = np.fromfile(tmp_path / "sim3_out.bin", np.int32)
    P = _lib.ProjParams()
    P.minX, P.maxX, P.minY, P.maxY = 0, w, 0, h
    P.grid_w_inv, P.grid_h_inv = float(kf.mfGridElementWidthInv), float(kf.mfGridElementHeightInv)
    P.nlevels = 8
    for i, sfv in enumerate(kf.mvScaleFactors[:8]):
        P.scale[i] = float(sfv)
    P.th, P.th_high = float(th), 50
    pts = dict(u=u, v=v, level=level, valid=valid, desc=desc)
    en, ekm, epm = oracle_search_kf_projection(P, pts, kf, taken)
    exp = np.where(ekm >= 0, ekm, np.where(taken != 0, -2, -1)).astype(np.int32)
    assert raw[0] == en and np.array_equal(raw[1:], exp), f"{(raw[1:] != exp).sum()} keypoints differ"
    assert en > 300 and 0.3 < valid.mean() < 0.95


def test_cpp_search_by_sim3_dropin(tmp_path):
    """The C++ template ORBmatcher::SearchBySim3 (LoopClosing::ComputeSim3) run for real on two synthetic keyframes seeing
    the same 3-D points: relative Sim3 algebra, both projection directions and the already-matched masks on the host
    (restated here in numpy with the same float / double steps), two device window searches, mutual agreement — against
    the oracle's per-direction search."""
    from oracle_lib import oracle_fuse_search
    from orbslam2_b200 import _lib
    from orbslam2_b200.extractor import KP_DTYPE as KPD
    import orbslam2_b200 as ob
    exe = _build_exe(tmp_path)
    rng = np.random.default_rng(123)
    f32, f64 = np.float32, np.float64
    w, h, th = 640, 480, f32(7.5)
    fx = fy = f32(520.0); cx, cy = f32(319.5), f32(240.25)
    sf = np.ones(8, f32)
    for i in range(1, 8):
        sf[i] = f32(f64(sf[i - 1]) * f64(f32(1.2)))
    n = 1400
    z = rng.uniform(2.5, 10, n)
    X = np.stack([(rng.uniform(30, w - 30, n) - cx) / fx * z, (rng.uniform(30, h - 30, n) - cy) / fy * z, z], 1)
    ang = 0.06
    R2 = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    t2 = np.array([0.15, -0.03, 0.1])
    R1w, t1w = np.eye(3, dtype=f32), np.zeros(3, f32)
    R2w, t2w = R2.astype(f32), t2.astype(f32)
    s12 = f32(1.02)
    R12 = (R1w.astype(f64) @ R2w.astype(f64).T).astype(f32)
    t12 = (t1w.astype(f64) - R12.astype(f64) @ t2w.astype(f64) + rng.normal(0, 0.01, 3)).astype(f32)
    base = rng.integers(0, 256, (n, 32), dtype=np.uint8)

    def flips(d, k):
        b = np.unpackbits(d, axis=1)
        for _ in range(k):
            b[np.arange(len(d)), rng.integers(0, 256, len(d))] ^= 1
        return np.packbits(b, axis=1)

    def make_side(Rw, tw, perm):
        pc = X[perm] @ Rw.astype(f64).T + tw.astype(f64)
        k = np.zeros(n, KPD)
        k["x"] = (fx * pc[:, 0] / pc[:, 2] + cx + rng.normal(0, 0.8, n)).astype(f32)
        k["y"] = (fy * pc[:, 1] / pc[:, 2] + cy + rng.normal(0, 0.8, n)).astype(f32)
        k["octave"] = rng.integers(0, 8, n); k["angle"] = rng.uniform(0, 360, n); k["class_id"] = -1
        k["size"] = 31 * sf[k["octave"]]
        return dict(keys=k, kd=flips(base[perm], 4), has=(rng.random(n) < 0.9).astype(np.uint8),
                    bad=(rng.random(n) < 0.04).astype(np.uint8), world=(X[perm] + rng.normal(0, 0.003, (n, 3))).astype(f32),
                    dmin=rng.uniform(0.5, 2.4, n).astype(f32), dmax=rng.uniform(8.0, 14.0, n).astype(f32), md=flips(base[perm], 5),
                    perm=perm)

    S1, S2 = make_side(R1w, t1w, np.arange(n)), make_side(R2w, t2w, rng.permutation(n))
    inv2 = np.empty(n, np.int64); inv2[S2["perm"]] = np.arange(n)
    # predicted level of a point in the OTHER keyframe: the octave of its true partner there (+0/1)
    S1["level"] = np.clip(S2["keys"]["octave"][inv2[S1["perm"]]] + rng.integers(0, 2, n), 0, 7).astype(np.int32)
    S2["level"] = np.clip(S1["keys"]["octave"][S2["perm"]] + rng.integers(0, 2, n), 0, 7).astype(np.int32)
    pre = np.where(rng.random(n) < 0.1, inv2[S1["perm"]], -1).astype(np.int32)   # matches found earlier by SearchByFboW
    with open(tmp_path / "sim3b.bin", "wb") as f:
        f.write(np.array([n, n], np.int32).tobytes())
        f.write(np.array([th, fx, fy, cx, cy, w, h, s12], f32).tobytes())
        for a in (R12, t12, R1w, t1w, R2w, t2w, sf):
            f.write(np.ascontiguousarray(a, f32).tobytes())
        for S in (S1, S2):
            f.write(S["keys"].tobytes())
            for key in ("kd", "has", "bad", "world", "dmin", "dmax", "level", "md"):
                f.write(np.ascontiguousarray(S[key]).tobytes())
        f.write(pre.tobytes())
    r = subprocess.run([exe, "sim3b", str(tmp_path / "sim3b.bin"), str(tmp_path / "sim3b_out.bin")], capture_output=True,
                       text=True, timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = np.fromfile(tmp_path / "sim3b_out.bin", np.int32)

    # numpy restatement of the template's host half (src/ORBmatcher.cc:1108-1121, 1140-1177, 1221-1258)
    def mul_add(A, x, b):
        return np.array([f32(f32(sum(f64(A[r, c]) * f64(x[c]) for c in range(3))) + b[r]) for r in range(3)], f32)
    inv = f32(1.0 / f64(s12))
    sR12 = (s12 * R12).astype(f32); sR21 = (inv * R12.T).astype(f32)
    t21 = np.array([-f32(sum(f64(sR21[r, c]) * f64(t12[c]) for c in range(3))) for r in range(3)], f32)
    already1 = pre >= 0
    already2 = np.zeros(n, bool); already2[pre[pre >= 0]] = True

    def direction(S, already, Rsw, tsw, sR, t, dstS):
        u = np.zeros(n, f32); v = np.zeros(n, f32); valid = np.zeros(n, np.uint8)
        for i in range(n):
            if not S["has"][i] or already[i] or S["bad"][i]:
                continue
            c2 = mul_add(sR, mul_add(Rsw, S["world"][i], tsw), t)
            if c2[2] < 0:
                continue
            invz = f32(1.0 / f64(c2[2]))
            uu = f32(f32(fx * f32(c2[0] * invz)) + cx); vv = f32(f32(fy * f32(c2[1] * invz)) + cy)
            if not (0 <= uu < w and 0 <= vv < h):
                continue
            d3 = f32(np.sqrt(sum(f64(c) * f64(c) for c in c2)))
            if d3 < S["dmin"][i] or d3 > S["dmax"][i]:
                continue
            u[i], v[i], valid[i] = uu, vv, 1
        kf = ob.Frame(dstS["keys"], dstS["kd"], sf, fx, fy, cx, cy, 0, w, h)
        P = _lib.ProjParams()
        P.minX, P.maxX, P.minY, P.maxY = 0, w, 0, h
        P.grid_w_inv, P.grid_h_inv = float(kf.mfGridElementWidthInv), float(kf.mfGridElementHeightInv)
        P.nlevels = 8
        for i, sfv in enumerate(sf):
            P.scale[i] = float(sfv)
        P.th = float(th)
        pts = dict(u=u, v=v, ur=np.zeros(n, f32), level=S["level"], valid=valid, desc=S["md"])
        bi, bd = oracle_fuse_search(P, np.zeros(16, f32), 0, pts, kf)
        return np.where((valid != 0) & (bi >= 0) & (bd <= 100), bi, -1), valid

    m1, v1 = direction(S1, already1, R1w, t1w, sR21, t21, S2)
    m2, v2 = direction(S2, already2, R2w, t2w, sR12, t12, S1)
    exp = np.where(pre >= 0, pre, -1).astype(np.int32)
    found = 0
    for i1 in range(n):
        if m1[i1] >= 0 and m2[m1[i1]] == i1:
            exp[i1] = m1[i1] if S2["has"][m1[i1]] else -1   # vpMatches12[i1] = vpMapPoints2[idx2] (NULL if KF2 has none there)
            found += 1
    assert raw[0] == found and np.array_equal(raw[1:], exp), f"{(raw[1:] != exp).sum()} entries differ"
    assert found > 300 and v1.mean() > 0.5 and v2.mean() > 0.5


def test_cpp_fuse_dropin(tmp_path):
    """The C++ template ORBmatcher::Fuse(KeyFrame*, vpMapPoints, th) (LocalMapping::SearchInNeighbors) run for real:
    projection and MapPoint gates on the host (restated in numpy with the same float / double steps), KeyFrame window
    search with the octave and chi-square gates on the device, Replace / AddObservation / AddMapPoint in point order."""
    from orbslam2_b200.synth import synth_tracking_pair
    from oracle_lib import oracle_fuse_search
    from orbslam2_b200 import _lib
    exe = _build_exe(tmp_path)
    rng = np.random.default_rng(321)
    f32, f64 = np.float32, np.float64
    kf, _ = synth_tracking_pair(1500, 55, stereo=True, motion=(0.03, 0.01, 0.05))
    w, h, th = 640, 480, f32(3.0)
    ang = -0.03
    R = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
    Rcw = R.astype(f32); tcw = np.array([0.05, 0.02, -0.1], f32)
    Ow = np.array([-f32(sum(f64(Rcw[c, r]) * f64(tcw[c]) for c in range(3))) for r in range(3)], f32)
    n = 2200
    src = rng.integers(0, kf.N, n)
    dupmask = rng.random(n) < 0.35
    src[dupmask] = src[rng.integers(0, 80, dupmask.sum())]          # several points aim at the same keypoint
    z = rng.uniform(2, 10, n)
    pc = np.stack([(kf.mvKeysUn["x"][src] + rng.normal(0, 0.7, n) - float(kf.cx)) / float(kf.fx) * z,
                   (kf.mvKeysUn["y"][src] + rng.normal(0, 0.7, n) - float(kf.cy)) / float(kf.fy) * z, z], 1)
    pc[rng.random(n) < 0.04, 2] *= -1
    world = ((pc - tcw.astype(f64)) @ R).astype(f32)
    bits = np.unpackbits(kf.mDescriptors[src], axis=1)
    for _ in range(8):
        bits[np.arange(n), rng.integers(0, 256, n)] ^= 1
    desc = np.packbits(bits, axis=1)
    level = np.clip(kf.mvKeysUn["octave"][src] + rng.integers(0, 2, n), 0, 7).astype(np.int32)
    bad = (rng.random(n) < 0.05).astype(np.uint8)
    nobs = rng.integers(1, 12, n).astype(np.int32)
    dmin = rng.uniform(0.5, 3.0, n).astype(f32); dmax = rng.uniform(6.0, 14.0, n).astype(f32)
    normal = np.zeros((n, 3), f32)
    for i in range(n):
        po = (world[i] - Ow).astype(f64)
        nv = po / np.linalg.norm(po)
        normal[i] = (-nv if rng.random() < 0.08 else nv).astype(f32)
    sf = np.asarray(kf.mvScaleFactors[:8], f32)
    is2 = (f32(1.0) / (sf * sf)).astype(f32)
    kfobs = np.where(rng.random(kf.N) < 0.6, rng.integers(0, 12, kf.N) + 1000 * (rng.random(kf.N) < 0.1), -1).astype(np.int32)
    # the template's host half (include/dropin/ORBmatcher.h Fuse) = src/ORBmatcher.cc:855-905
    u = np.zeros(n, f32); v = np.zeros(n, f32); ur = np.zeros(n, f32); valid = np.zeros(n, np.uint8)
    for i in range(n):
        if bad[i]:
            continue
        wv = world[i]
        p = [f32(f32(sum(f64(Rcw[r, c]) * f64(wv[c]) for c in range(3))) + tcw[r]) for r in range(3)]
        if p[2] < 0:
            continue
        invz = f32(f32(1) / p[2])
        uu = f32(f32(kf.fx * f32(p[0] * invz)) + kf.cx); vv = f32(f32(kf.fy * f32(p[1] * invz)) + kf.cy)
        if not (0 <= uu < w and 0 <= vv < h):
            continue
        po = [f32(wv[k] - Ow[k]) for k in range(3)]
        dist = f32(np.sqrt(sum(f64(x) * f64(x) for x in po)))
        if dist < dmin[i] or dist > dmax[i]:
            continue
        if sum(f64(po[k]) * f64(normal[i, k]) for k in range(3)) < 0.5 * f64(dist):
            continue
        u[i], v[i], ur[i], valid[i] = uu, vv, f32(uu - f32(kf.mbf * invz)), 1
    with open(tmp_path / "fuse.bin", "wb") as f:
        f.write(np.array([n, kf.N], np.int32).tobytes())
        f.write(np.array([th, kf.fx, kf.fy, kf.cx, kf.cy, kf.mbf, w, h], f32).tobytes())
        f.write(Rcw.tobytes()); f.write(tcw.tobytes()); f.write(Ow.tobytes()); f.write(sf.tobytes()); f.write(is2.tobytes())
        for a in (world, normal, dmin, dmax, level, nobs, bad, desc):
            f.write(np.ascontiguousarray(a).tobytes())
        f.write(kf.mvKeysUn.tobytes()); f.write(np.ascontiguousarray(kf.mvuRight, f32).tobytes())
        f.write(kf.mDescriptors.tobytes()); f.write(kfobs.tobytes())
    r = subprocess.run([exe, "fuse", str(tmp_path / "fuse.bin"), str(tmp_path / "fuse_out.bin")], capture_output=True, text=True,
                       timeout=120)
    assert r.returncode == 0, r.stdout + r.stderr
    raw = np.fromfile(tmp_path / "fuse_out.bin", np.int32)
    P = _lib.ProjParams()
    P.minX, P.maxX, P.minY, P.maxY = 0, w, 0, h
    P.grid_w_inv, P.grid_h_inv = float(kf.mfGridElementWidthInv), float(kf.mfGridElementHeightInv)
    P.nlevels = 8
    for i, sfv in enumerate(sf):
        P.scale[i] = float(sfv)
    P.th = float(th)
    inv16 = np.zeros(16, f32); inv16[:8] = is2
    bi, bd = oracle_fuse_search(P, inv16, 1, dict(u=u, v=v, ur=ur, level=level, valid=valid, desc=desc), kf)
    # the mutation loop in point order (src/ORBmatcher.cc:948-967)
    action = np.zeros(n, np.int32); where = np.full(n, -1, np.int32)
    holder = {i: ("own", i) for i in range(kf.N) if kfobs[i] >= 0}
    own_replaced = set()
    fused = 0
    for i in range(n):
        if not valid[i] or bad[i] or bi[i] < 0 or bd[i] > 50:
            continue
        idx = int(bi[i])
        if idx in holder:
            kind, j = holder[idx]
            hbad = (kfobs[j] >= 1000) if kind == "own" else bool(bad[j])
            hobs = (kfobs[j] % 1000) if kind == "own" else nobs[j]
            if not hbad:
                if hobs > nobs[i]:
                    action[i] = -1                         # pMP->Replace(pMPinKF)
                elif kind == "own":
                    own_replaced.add(j)                    # pMPinKF->Replace(pMP)
                else:
                    action[j] = -1
        else:
            action[i], where[i] = 3, idx                   # AddObservation + AddMapPoint
            holder[idx] = ("pt", i)
        fused += 1
    assert raw[0] == fused, (raw[0], fused)
    got = raw[1:1 + 2 * n].reshape(n, 2)
    assert np.array_equal(got[:, 0], action) and np.array_equal(got[:, 1], where)
    assert raw[1 + 2 * n] == len(own_replaced)
    assert fused > 300 and (action == 3).sum() > 50 and (action == -1).sum() > 20 and len(own_replaced) > 20
