"""Scenes for tests/cpp/scene_run.cpp — TEST INFRASTRUCTURE.

A scene is a small synthetic SLAM world (views = frames / keyframes with keypoints, descriptors, poses, feature vectors
and map-point associations; map points with positions, normals, descriptors, observations) plus the arguments of ONE
ORBmatcher method. scene_run executes that method and dumps the return value, the output vectors and the whole map
state. The same scene file goes through
  * oracle/_ref/scene_run_ref          the UNMODIFIED /root/reference/src/ORBmatcher.cc (the pin),
  * scene_run over liborbx.so          the drop-in templates on the GPU (tests/test_gpu_refmatcher.py),
  * scene_run over the oracle backend  drop-in host halves + orc_* on the CPU (tests/test_cpu_refmatcher.py),
and the dumps must be identical int32 streams.
"""
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
REF_EXE = os.path.join(REF_DIR, "scene_run_ref")
REF_SO = os.path.join(REF_DIR, "libref_orbmatcher.so")
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
f32, f64 = np.float32, np.float64

LOCAL, FRAME, RELOC, KFPROJ, BOW_KF_F, BOW_KF_KF, INIT, TRIANG, SIM3, FUSE, FUSE_SIM3 = range(1, 12)
METHOD_NAMES = {LOCAL: "SearchByProjection(F,vpMapPoints)", FRAME: "SearchByProjection(F,F)",
                RELOC: "SearchByProjection(F,KF,found)", KFPROJ: "SearchByProjection(KF,Scw)",
                BOW_KF_F: "SearchByFboW(KF,F)", BOW_KF_KF: "SearchByFboW(KF,KF)", INIT: "SearchForInitialization",
                TRIANG: "SearchForTriangulation", SIM3: "SearchBySim3", FUSE: "Fuse(KF,vpMapPoints)",
                FUSE_SIM3: "Fuse(KF,Scw)"}


def have_ref_matcher():
    return os.path.exists(REF_EXE) and os.path.exists(REF_SO)


def scale_tables(nlevels=8, scale=1.2):
    """mvScaleFactor / mvLevelSigma2 and inverses as the ORBextractor constructor fills them
    (/root/reference/src/ORBextractor.cc:417-432): float products, float reciprocals."""
    sf = np.ones(nlevels, f32); s2 = np.ones(nlevels, f32)
    for i in range(1, nlevels):
        sf[i] = f32(sf[i - 1] * f32(scale))
        s2[i] = f32(sf[i] * sf[i])
    return sf, (f32(1.0) / sf).astype(f32), s2, (f32(1.0) / s2).astype(f32)


class View:
    """One frame or keyframe of a scene."""

    def __init__(self, keys, desc, u_right, depth, Tcw, kind=0, keys_un=None):
        self.kind = kind
        self.keys = np.ascontiguousarray(keys, KP_DTYPE)
        self.keys_un = self.keys.copy() if keys_un is None else np.ascontiguousarray(keys_un, KP_DTYPE)
        self.N = len(self.keys)
        self.desc = np.ascontiguousarray(desc, np.uint8).reshape(self.N, 32)
        self.u_right = np.ascontiguousarray(u_right, f32)
        self.depth = np.ascontiguousarray(depth, f32)
        self.Tcw = np.ascontiguousarray(Tcw, f32).reshape(4, 4)
        self.featvec = {}                                  # node id -> ascending feature indices
        self.mp = np.full(self.N, -1, np.int32)            # mvpMapPoints as map-point ids
        self.outlier = np.zeros(self.N, np.uint8)
        self.phys = np.full(self.N, -1, np.int32)          # generator bookkeeping: physical point behind a keypoint


class MapPointRec:
    def __init__(self, world, normal, desc, min_dist, max_dist):
        self.world = np.asarray(world, f32); self.normal = np.asarray(normal, f32)
        self.desc = np.asarray(desc, np.uint8)
        self.min_dist, self.max_dist = f32(min_dist), f32(max_dist)
        self.bad = 0; self.nobs_override = -1; self.visible = 1; self.found = 1
        self.in_view = 0; self.track_level = 0
        self.proj = np.zeros(4, f32)                       # mTrackProjX, mTrackProjY, mTrackProjXR, mTrackViewCos
        self.obs = []                                      # (view id, keypoint idx); the view must be a keyframe


class Scene:
    def __init__(self, method, cal, nlevels=8, scale=1.2):
        self.method = method
        self.cal = dict(cal)
        self.nlevels, self.scale = nlevels, f32(scale)
        self.sf, self.isf, self.s2, self.is2 = scale_tables(nlevels, scale)
        self.views, self.mps = [], []
        self.fpar = np.zeros(8, f32)    # [0] th  [1] nnratio  [2] s12  [4..7] fx, fy, cx, cy of keyframe view 1 (0 = shared)
        self.ipar = np.zeros(8, np.int32)  # [0] checkOrientation  [1] method-specific int
        self.lists, self.farr = [], []

    def tobytes(self):
        c = self.cal
        out = [np.array([0x53434E31, self.method, len(self.views), len(self.mps)], np.int32).tobytes(),
               self.fpar.tobytes(), self.ipar.tobytes(),
               np.array([c["fx"], c["fy"], c["cx"], c["cy"], c["bf"], c["b"], c["th_depth"], c["min_x"], c["max_x"],
                         c["min_y"], c["max_y"], f32(np.log(f64(self.scale)))], f32).tobytes(),  # mfLogScaleFactor = log(mfScaleFactor)
               np.array([self.nlevels], np.int32).tobytes(), np.array([self.scale], f32).tobytes(),
               self.sf.tobytes(), self.isf.tobytes(), self.s2.tobytes(), self.is2.tobytes()]
        for v in self.views:
            out += [np.array([v.kind, v.N], np.int32).tobytes(), v.keys.tobytes(), v.keys_un.tobytes(), v.u_right.tobytes(),
                    v.depth.tobytes(), v.desc.tobytes(), v.Tcw.tobytes()]
            nodes = sorted(v.featvec)
            off = np.zeros(len(nodes) + 1, np.int32)
            for j, k in enumerate(nodes):
                off[j + 1] = off[j] + len(v.featvec[k])
            feats = np.concatenate([np.asarray(v.featvec[k], np.uint32) for k in nodes]) if nodes else np.zeros(0, np.uint32)
            out += [np.array([len(nodes)], np.int32).tobytes(), np.asarray(nodes, np.uint32).tobytes(), off.tobytes(),
                    feats.astype(np.uint32).tobytes(), v.mp.astype(np.int32).tobytes(), v.outlier.astype(np.uint8).tobytes()]
        for m in self.mps:
            out += [np.concatenate([m.world, m.normal, [m.min_dist, m.max_dist]]).astype(f32).tobytes(), m.desc.tobytes(),
                    np.array([m.bad, m.nobs_override, m.visible, m.found, m.in_view, m.track_level], np.int32).tobytes(),
                    m.proj.astype(f32).tobytes(), np.array([len(m.obs)], np.int32).tobytes(),
                    np.asarray(m.obs, np.int32).reshape(-1, 2).tobytes()]
        out.append(np.array([len(self.lists)], np.int32).tobytes())
        for l in self.lists:
            l = np.asarray(l, np.int32)
            out += [np.array([len(l)], np.int32).tobytes(), l.tobytes()]
        out.append(np.array([len(self.farr)], np.int32).tobytes())
        for a in self.farr:
            a = np.asarray(a, f32).ravel()
            out += [np.array([len(a)], np.int32).tobytes(), a.tobytes()]
        return b"".join(out)


def run_scene(exe, scene, tmp, tag):
    """Run one scene_run variant; returns the int32 dump."""
    sp, op = os.path.join(str(tmp), f"{tag}.scene"), os.path.join(str(tmp), f"{tag}.out")
    with open(sp, "wb") as f:
        f.write(scene.tobytes())
    r = subprocess.run([exe, sp, op], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, f"{exe}: rc {r.returncode}\n{r.stdout}\n{r.stderr}"
    return np.fromfile(op, np.int32)


def parse_dump(scene, d):
    """Split a dump into {ret, out: [lists], views: [mvpMapPoints ids], mps: [dict]}."""
    at = 2 + int(d[1])
    res = {"ret": int(d[0]), "out": [], "views": [], "mps": []}
    o = d[2:at]; k = 0
    while k < len(o):
        n = int(o[k]); res["out"].append(o[k + 1:k + 1 + n]); k += 1 + n
    for _ in scene.views:
        n = int(d[at]); res["views"].append(d[at + 1:at + 1 + n]); at += 1 + n
    for _ in scene.mps:
        nob = int(d[at + 5])
        res["mps"].append({"bad": int(d[at]), "nObs": int(d[at + 1]), "replaced": int(d[at + 2]), "visible": int(d[at + 3]),
                           "found": int(d[at + 4]), "obs": d[at + 6:at + 6 + 2 * nob].reshape(-1, 2),
                           "desc": d[at + 6 + 2 * nob:at + 14 + 2 * nob].copy().view(np.uint8)})
        at += 14 + 2 * nob
    assert at == len(d)
    return res


def _compile(out, extra_src, libs, tmp):
    exe = os.path.join(str(tmp), out)
    cmd = ["g++", "-std=c++14", "-O1", "-w", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "oracle", "shim_matcher"),
           os.path.join(ROOT, "tests", "cpp", "scene_run.cpp")] + extra_src + ["-o", exe]
    for d, l in libs:
        cmd += ["-L", d, "-l" + l, "-Wl,-rpath," + d]
    subprocess.check_call(cmd)
    return exe


def build_hostcheck(tmp):
    """scene_run with the drop-in templates over tests/cpp/orbm_oracle_backend.cpp (CPU)."""
    from oracle_lib import build_oracle
    build_oracle()
    return _compile("scene_run_host", [os.path.join(ROOT, "tests", "cpp", "orbm_oracle_backend.cpp")],
                    [(os.path.join(ROOT, "oracle"), "orb_oracle"), (REF_DIR, "ref_orbmatcher")], tmp)


def build_gpu(tmp):
    """scene_run with the drop-in templates over liborbx.so (the product)."""
    from orbslam2_b200 import build
    lib = build.build()
    return _compile("scene_run_gpu", [], [(os.path.dirname(lib), "orbx"), (REF_DIR, "ref_orbmatcher")], tmp)


# ------------------------------------------------------------------------------------------------------------------
# A synthetic world: physical 3-D points seen from several nearby poses.
# ------------------------------------------------------------------------------------------------------------------
def default_cal(width=640, height=480, stereo=True):
    fx = f32(520.0); bf = f32(40.0)
    return dict(fx=fx, fy=f32(520.0), cx=f32(width / 2 - 0.5), cy=f32(height / 2 + 0.25), bf=bf, b=f32(bf / fx),
                th_depth=f32(35.0 * bf / fx), min_x=f32(0), max_x=f32(width), min_y=f32(0), max_y=f32(height),
                width=width, height=height, stereo=stereo)


def pose(ry=0.0, rx=0.0, rz=0.0, t=(0, 0, 0)):
    cy_, sy = np.cos(ry), np.sin(ry); cx_, sx = np.cos(rx), np.sin(rx); cz, sz = np.cos(rz), np.sin(rz)
    Ry = np.array([[cy_, 0, sy], [0, 1, 0], [-sy, 0, cy_]]); Rx = np.array([[1, 0, 0], [0, cx_, -sx], [0, sx, cx_]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    T = np.eye(4); T[:3, :3] = Rz @ Ry @ Rx; T[:3, 3] = t
    return T.astype(f32)


class PhysWorld:
    """n physical points; view(k) observes them from poses[k]. Keypoints of the same physical point share a descriptor up
    to bit flips, a consistent orientation offset and a distance-consistent octave."""

    def __init__(self, seed, n_points, poses, cal, nlevels=8, scale=1.2, vocab_blob=None, dup_frac=0.1, knife=None):
        self.rng = rng = np.random.default_rng(seed)
        self.cal, self.nlevels = cal, nlevels
        self.sf, self.isf, self.s2, self.is2 = scale_tables(nlevels, scale)
        self.poses = [np.asarray(p, f32) for p in poses]
        w, h = cal["width"], cal["height"]
        z = rng.uniform(2.0, 12.0, n_points)
        u = rng.uniform(-40, w + 40, n_points); v = rng.uniform(-30, h + 30, n_points)
        self.X = np.stack([(u - cal["cx"]) / cal["fx"] * z, (v - cal["cy"]) / cal["fy"] * z, z], 1).astype(f32)
        if vocab_blob is None:
            self.desc = rng.integers(0, 256, (n_points, 32), dtype=np.uint8)
        else:
            from orbslam2_b200.synth import synth_vocabulary_features
            self.desc = synth_vocabulary_features(vocab_blob, n_points, seed=seed + 1, flips=6, random_frac=0.05)
        # near-duplicate physical points (repeated texture): same place, nearly the same descriptor
        nd = int(n_points * dup_frac)
        if nd:
            dup = rng.choice(n_points, nd, replace=False)
            src = rng.choice(np.setdiff1d(np.arange(n_points), dup), nd)
            self.X[dup] = self.X[src] + rng.normal(0, 0.004, (nd, 3)).astype(f32)
            self.desc[dup] = self._flip(self.desc[src], 2)
        self.n = n_points
        self.angle0 = rng.uniform(0, 360, n_points)
        self.level0 = rng.integers(0, nlevels - 1, n_points)      # octave at the reference distance
        # reference distance of a point = its distance to one of the `knife` views (default: the origin). A search from
        # that view then evaluates PredictScale at log(ratio)/log(1.2) = an integer +- a few ulp, so that a last-bit
        # difference in the projected point (float vs double accumulation, a fused multiply-add) changes the octave gate
        # and shows up in the matches: the scenes are deliberately ill-conditioned.
        if knife is None:
            d0 = np.linalg.norm(self.X.astype(f64), axis=1)
        else:
            cs = np.stack([self.centre(k) for k in knife])
            d0 = np.linalg.norm(self.X.astype(f64) - cs[rng.integers(0, len(knife), n_points)], axis=1)
        self.max_dist = (d0 * self.sf[self.level0]).astype(f32)   # MapPoint::UpdateNormalAndDepth, src/MapPoint.cc:363-366
        self.min_dist = (self.max_dist / self.sf[nlevels - 1]).astype(f32)
        self.normal = (self.X.astype(f64) / d0[:, None]).astype(f32)

    def _flip(self, desc, nbits):
        bits = np.unpackbits(np.atleast_2d(desc), axis=1)
        n = len(bits)
        nb = np.broadcast_to(nbits, (n,))
        for k in range(int(nb.max()) if n else 0):
            sel = np.flatnonzero(nb > k)
            bits[sel, self.rng.integers(0, 256, len(sel))] ^= 1
        return np.packbits(bits, axis=1)

    def project(self, k, X=None):
        T = self.poses[k].astype(f64)
        Xc = (T[:3, :3] @ (self.X if X is None else X).astype(f64).T).T + T[:3, 3]
        c = self.cal
        with np.errstate(divide="ignore", invalid="ignore"):
            u = c["fx"] * Xc[:, 0] / Xc[:, 2] + c["cx"]; v = c["fy"] * Xc[:, 1] / Xc[:, 2] + c["cy"]
        return u, v, Xc[:, 2]

    def centre(self, k):
        T = self.poses[k].astype(f64)
        return -T[:3, :3].T @ T[:3, 3]

    def predict_level(self, k, idx=None):
        idx = np.arange(self.n) if idx is None else idx
        dist = np.linalg.norm(self.X[idx].astype(f64) - self.centre(k), axis=1)
        ratio = self.max_dist[idx] / dist
        lvl = np.ceil(np.log(ratio) / np.log(1.2)).astype(int)
        return np.clip(lvl, 0, self.nlevels - 1), dist

    def view(self, k, kind, detect=0.85, noise=1.2, flip_p=0.05, clutter=0.3, wild=0.08, mono_frac=0.2, roll=0.0,
             margin=2.0, level_jitter=True):
        rng, c = self.rng, self.cal
        u, v, z = self.project(k)
        ok = (z > 0.5) & (u > margin) & (u < c["width"] - margin) & (v > margin) & (v < c["height"] - margin)
        ok &= rng.random(self.n) < detect
        ids = np.flatnonzero(ok)
        n = len(ids); nc = int(n * clutter)
        x = np.concatenate([u[ids] + rng.normal(0, noise, n), rng.uniform(1, c["width"] - 1, nc)]).astype(f32)
        y = np.concatenate([v[ids] + rng.normal(0, noise, n), rng.uniform(1, c["height"] - 1, nc)]).astype(f32)
        x = np.clip(x, 0.5, c["width"] - 1.5).astype(f32); y = np.clip(y, 0.5, c["height"] - 1.5).astype(f32)
        lvl, _ = self.predict_level(k, ids)
        if level_jitter:
            lvl = np.clip(lvl - (rng.random(n) < 0.5) + (rng.random(n) < 0.1), 0, self.nlevels - 1)
        octv = np.concatenate([lvl, rng.integers(0, self.nlevels, nc)]).astype(np.int32)
        ang = np.concatenate([self.angle0[ids] + roll + rng.normal(0, 3, n), rng.uniform(0, 360, nc)])
        w = np.flatnonzero(rng.random(n) < wild)
        ang[w] = rng.uniform(0, 360, len(w))
        ang = np.mod(ang, 360).astype(f32)
        desc = np.concatenate([self._flip(self.desc[ids], rng.binomial(256, flip_p, n)),
                               rng.integers(0, 256, (nc, 32), dtype=np.uint8)])
        zz = np.concatenate([z[ids], rng.uniform(2, 12, nc)])
        ur = (x - c["bf"] / zz).astype(f32); dp = zz.astype(f32)
        nomatch = rng.random(n + nc) < (mono_frac if c["stereo"] else 1.1)
        ur[nomatch] = -1; dp[nomatch] = -1
        phys = np.concatenate([ids, np.full(nc, -1)]).astype(np.int32)
        perm = rng.permutation(n + nc)
        keys = np.zeros(n + nc, KP_DTYPE)
        keys["x"], keys["y"], keys["octave"], keys["angle"] = x[perm], y[perm], octv[perm], ang[perm]
        keys["size"] = 31 * self.sf[octv[perm]]; keys["class_id"] = -1
        V = View(keys, desc[perm], ur[perm], dp[perm], self.poses[k], kind)
        V.phys = phys[perm]
        return V

    def map_point(self, j, desc=None):
        return MapPointRec(self.X[j], self.normal[j], self.desc[j] if desc is None else desc, self.min_dist[j], self.max_dist[j])


def attach_featvec(view, voc, level=4):
    """mFbowFeatVec = Vocabulary::transform(mDescriptors, 4, ...) through the oracle's fbow restatement (itself pinned
    to the unmodified fbow.cpp, tests/test_cpu_fbow.py)."""
    _, fv = voc.transform(view.desc, level)
    view.featvec = {int(k): np.asarray(v, np.uint32) for k, v in fv.items()}


# ------------------------------------------------------------------------------------------------------------------
# Scene builders, one per ORBmatcher method. Each returns a Scene whose arguments mirror the reference's call site.
# ------------------------------------------------------------------------------------------------------------------
def _new_scene(method, world, th, nnratio, check_ori=True, ip1=0):
    S = Scene(method, world.cal, world.nlevels)
    S.fpar[0], S.fpar[1] = th, nnratio
    S.ipar[0], S.ipar[1] = int(check_ori), ip1
    return S


def _mps_for_view(S, world, V, frac=0.95, flips=3):
    """One map point per keypoint of V that has a physical point (probability frac); returns keypoint -> mp id."""
    rng = world.rng
    for i in range(V.N):
        j = V.phys[i]
        if j >= 0 and rng.random() < frac:
            V.mp[i] = len(S.mps)
            S.mps.append(world.map_point(j, world._flip(world.desc[j], rng.integers(0, flips + 1))[0]))


def scene_frame(seed, stereo=True, bMono=False, th=15.0, n_points=1500, motion=(0.05, -0.02, 0.12), check_ori=True):
    """Tracking::TrackWithMotionModel -> SearchByProjection(CurrentFrame, LastFrame, th, bMono), src/Tracking.cc:985."""
    cal = default_cal(stereo=stereo)
    W = PhysWorld(seed, n_points, [pose(0.02, t=motion), pose()], cal)
    S = _new_scene(FRAME, W, th, 0.9, check_ori, int(bMono))
    cur = W.view(0, 0); last = W.view(1, 0, clutter=0.1)
    _mps_for_view(S, W, last)
    rng = W.rng
    last.outlier[rng.random(last.N) < 0.03] = 1
    for m in S.mps:                                  # temporal (visual-odometry) points have no observations
        m.nobs_override = 0 if rng.random() < 0.15 else int(rng.integers(1, 6))
    # some current keypoints are already occupied (by observed and by unobserved points)
    for i in np.flatnonzero(rng.random(cur.N) < 0.06):
        cur.mp[i] = len(S.mps)
        m = W.map_point(int(rng.integers(0, W.n))); m.nobs_override = int(rng.integers(0, 3))
        S.mps.append(m)
    S.views = [cur, last]
    return S


def scene_local(seed, stereo=True, th=3.0, nnratio=0.8, n_points=1500):
    """Tracking::SearchLocalPoints -> SearchByProjection(mCurrentFrame, mvpLocalMapPoints, th), src/Tracking.cc:1290.
    The mTrack* members are what Frame::isInFrustum (src/Frame.cc:270-326) left in the points."""
    cal = default_cal(stereo=stereo)
    W = PhysWorld(seed, n_points, [pose(0.015, t=(0.04, 0.01, 0.08))], cal)
    S = _new_scene(LOCAL, W, th, nnratio)
    F = W.view(0, 0)
    rng = W.rng
    u, v, z = W.project(0)
    lvl, dist = W.predict_level(0)
    ids = []
    for j in range(W.n):
        m = W.map_point(j)
        inside = z[j] > 0 and 0 <= u[j] < cal["width"] and 0 <= v[j] < cal["height"]
        m.in_view = int(inside and rng.random() < 0.9)
        m.bad = int(rng.random() < 0.03)
        m.nobs_override = int(rng.integers(1, 8))
        if inside:
            m.proj[:] = [u[j] + rng.normal(0, 0.8), v[j] + rng.normal(0, 0.8), u[j] - cal["bf"] / z[j], rng.uniform(0.5, 1.0)]
            m.track_level = int(lvl[j])
        ids.append(len(S.mps)); S.mps.append(m)
    # the frame already tracks some points (TrackWithMotionModel ran first): those block their keypoints
    for i in np.flatnonzero(rng.random(F.N) < 0.25):
        if F.phys[i] >= 0:
            F.mp[i] = len(S.mps)
            m = W.map_point(int(F.phys[i])); m.nobs_override = int(rng.integers(0, 4))
            S.mps.append(m)
    S.views = [F]
    S.lists = [rng.permutation(ids)]
    return S


def scene_reloc(seed, th=10.0, orb_dist=100, n_points=1500):
    """Tracking::Relocalization -> SearchByProjection(mCurrentFrame, vpCandidateKFs[i], sFound, 10, 100),
    src/Tracking.cc:1552 (and :1566 with th 3, ORBdist 64)."""
    cal = default_cal(stereo=True)
    W = PhysWorld(seed, n_points, [pose(0.03, 0.01, t=(0.08, -0.03, 0.15)), pose()], cal, knife=[0, 1])
    S = _new_scene(RELOC, W, th, 0.9, True, orb_dist)
    cur = W.view(0, 0); kf = W.view(1, 1, clutter=0.1)
    _mps_for_view(S, W, kf)
    rng = W.rng
    for i in np.flatnonzero(kf.mp >= 0):
        S.mps[kf.mp[i]].obs.append((1, int(i)))
        S.mps[kf.mp[i]].bad = int(rng.random() < 0.03)
    # points found by the earlier PnP inliers: in the frame already and in sAlreadyFound
    found = []
    phys_to_mp = {int(kf.phys[i]): int(kf.mp[i]) for i in np.flatnonzero(kf.mp >= 0)}
    for i in range(cur.N):
        j = int(cur.phys[i])
        if j in phys_to_mp and rng.random() < 0.25:
            cur.mp[i] = phys_to_mp[j]; found.append(phys_to_mp[j])
    S.views = [cur, kf]
    S.lists = [found]
    return S


def _sim3_of(T, s):
    """Scw = [s*R | s*t ... ] as LoopClosing builds it from a g2o::Sim3 (src/LoopClosing.cc:394-397: sR, t)."""
    Scw = np.eye(4, dtype=f32)
    Scw[:3, :3] = (f64(s) * T[:3, :3].astype(f64)).astype(f32)
    Scw[:3, 3] = T[:3, 3]
    return Scw


def scene_kfproj(seed, th=10, s=1.0, n_points=1500, dup=0.3):
    """LoopClosing::ComputeSim3 -> SearchByProjection(mpCurrentKF, mScw, mvpLoopMapPoints, mvpCurrentMatchedPoints, 10),
    src/LoopClosing.cc:411."""
    cal = default_cal(stereo=True)
    W = PhysWorld(seed, n_points, [pose(-0.02, 0.01, t=(0.05, 0.02, -0.1))], cal, dup_frac=dup, knife=[0])
    S = _new_scene(KFPROJ, W, 0, 0.75, True, th)
    kf = W.view(0, 1)
    rng = W.rng
    ids = []
    for j in range(W.n):
        m = W.map_point(j, W._flip(W.desc[j], rng.integers(0, 6))[0])
        m.bad = int(rng.random() < 0.04)
        if rng.random() < 0.06:
            m.normal = (-m.normal).astype(f32)             # fails the viewing-angle gate
        m.nobs_override = int(rng.integers(1, 9))
        ids.append(len(S.mps)); S.mps.append(m)
    # vpMatched: what SearchByFboW + the Sim3 inliers already fixed, indexed by keypoint of the keyframe
    matched = np.full(kf.N, -1, np.int32)
    for i in np.flatnonzero((kf.phys >= 0) & (rng.random(kf.N) < 0.2)):
        matched[i] = ids[int(kf.phys[i])]
    S.views = [kf]
    S.lists = [rng.permutation(ids)[:int(0.8 * len(ids))], matched]
    # the Sim3 maps world -> camera with scale s: points are at s * (R X + t) ... keep them in front of the camera
    T = W.poses[0].copy()
    Scw = np.eye(4, dtype=f32)
    Scw[:3, :3] = (f64(s) * T[:3, :3].astype(f64)).astype(f32); Scw[:3, 3] = (f64(s) * T[:3, 3].astype(f64)).astype(f32)
    S.farr = [Scw]
    return S


def scene_init(seed, window=100, nnratio=0.9, check_ori=True, n_points=1500):
    """Tracking::MonocularInitialization -> SearchForInitialization(mInitialFrame, mCurrentFrame, mvbPrevMatched,
    mvIniMatches, 100), src/Tracking.cc:699."""
    cal = default_cal(stereo=False)
    W = PhysWorld(seed, n_points, [pose(), pose(0.01, t=(0.06, 0.0, 0.02))], cal, dup_frac=0.25)
    S = _new_scene(INIT, W, 0, nnratio, check_ori, window)
    F1 = W.view(0, 0, level_jitter=False); F2 = W.view(1, 0)
    F1.keys["octave"][W.rng.random(F1.N) < 0.6] = 0; F1.keys_un = F1.keys.copy()   # only level-0 keypoints query (:417)
    F2.keys["octave"][W.rng.random(F2.N) < 0.6] = 0; F2.keys_un = F2.keys.copy()
    S.views = [F1, F2]
    prev = np.stack([F1.keys_un["x"], F1.keys_un["y"]], 1).astype(f32)
    S.farr = [prev]
    return S


_VOC = {}


def _vocab(k=10, levels=5, seed=21):
    key = (k, levels, seed)
    if key not in _VOC:
        from orbslam2_b200.synth import synth_vocabulary
        from oracle_lib import OracleVocabulary
        blob = synth_vocabulary(k=k, levels=levels, seed=seed)
        _VOC[key] = (blob, OracleVocabulary(blob))
    return _VOC[key]


def _observe(S, V, vid):
    for i in np.flatnonzero(V.mp >= 0):
        S.mps[V.mp[i]].obs.append((vid, int(i)))


def scene_bow_kf_f(seed, nnratio=0.7, n_points=1200, check_ori=True):
    """Tracking::TrackReferenceKeyFrame -> SearchByFboW(mpReferenceKF, mCurrentFrame, vpMapPointMatches),
    src/Tracking.cc:870 (ORBmatcher(0.7, true)); Relocalization uses (0.75, true), src/Tracking.cc:1465."""
    blob, voc = _vocab()
    cal = default_cal(stereo=True)
    W = PhysWorld(seed, n_points, [pose(), pose(0.03, t=(0.08, 0.0, 0.1), rz=0.05)], cal, vocab_blob=blob, dup_frac=0.2)
    S = _new_scene(BOW_KF_F, W, 0, nnratio, check_ori)
    kf = W.view(0, 1); F = W.view(1, 0, roll=3.0)
    _mps_for_view(S, W, kf, frac=0.8)
    _observe(S, kf, 0)
    for m in S.mps:
        m.bad = int(W.rng.random() < 0.04)
    attach_featvec(kf, voc); attach_featvec(F, voc)
    S.views = [kf, F]
    return S


def scene_bow_kf_kf(seed, nnratio=0.75, n_points=1200):
    """LoopClosing::ComputeSim3 -> SearchByFboW(mpCurrentKF, pKF, vvpMapPointMatches[i]), src/LoopClosing.cc:301."""
    blob, voc = _vocab()
    cal = default_cal(stereo=True)
    W = PhysWorld(seed, n_points, [pose(), pose(-0.04, t=(-0.1, 0.02, 0.05))], cal, vocab_blob=blob, dup_frac=0.2)
    S = _new_scene(BOW_KF_KF, W, 0, nnratio, True)
    k1 = W.view(0, 1); k2 = W.view(1, 1, roll=-4.0)
    _mps_for_view(S, W, k1, frac=0.8); _observe(S, k1, 0)
    n1 = len(S.mps)
    _mps_for_view(S, W, k2, frac=0.7)
    for i in np.flatnonzero(k2.mp >= 0):
        S.mps[k2.mp[i]].obs.append((1, int(i)))
    for m in S.mps:
        m.bad = int(W.rng.random() < 0.04)
    assert len(S.mps) > n1
    attach_featvec(k1, voc); attach_featvec(k2, voc)
    S.views = [k1, k2]
    return S


def _skew(t):
    return np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]], f64)


def scene_triang(seed, only_stereo=False, stereo=True, n_points=1200, check_ori=False):
    """LocalMapping::CreateNewMapPoints -> SearchForTriangulation(mpCurrentKeyFrame, pKF2, F12, vMatchedIndices, false),
    src/LocalMapping.cc:268 with F12 from ComputeF12 (src/LocalMapping.cc:550-568); ORBmatcher(0.6, false)."""
    blob, voc = _vocab()
    cal = default_cal(stereo=stereo)
    W = PhysWorld(seed, n_points, [pose(), pose(0.02, t=(-0.25, 0.01, 0.03))], cal, vocab_blob=blob, dup_frac=0.15)
    S = _new_scene(TRIANG, W, 0, 0.6, check_ori, int(only_stereo))
    k1 = W.view(0, 1, mono_frac=0.4); k2 = W.view(1, 1, mono_frac=0.4)
    # most keypoints have no map point yet (that is what gets triangulated); some do and are skipped
    _mps_for_view(S, W, k1, frac=0.3); _observe(S, k1, 0)
    _mps_for_view(S, W, k2, frac=0.3)
    for i in np.flatnonzero(k2.mp >= 0):
        S.mps[k2.mp[i]].obs.append((1, int(i)))
    attach_featvec(k1, voc); attach_featvec(k2, voc)
    S.views = [k1, k2]
    T1, T2 = W.poses[0].astype(f64), W.poses[1].astype(f64)
    R12 = T1[:3, :3] @ T2[:3, :3].T
    t12 = -R12 @ T2[:3, 3] + T1[:3, 3]
    K = np.array([[cal["fx"], 0, cal["cx"]], [0, cal["fy"], cal["cy"]], [0, 0, 1]], f64)
    F12 = np.linalg.inv(K).T @ _skew(t12) @ R12 @ np.linalg.inv(K)
    S.farr = [F12.astype(f32)]
    return S


def scene_sim3(seed, th=7.5, s12=1.0, n_points=1200, other_cal=False):
    """LoopClosing::ComputeSim3 -> SearchBySim3(mpCurrentKF, pKF, vpMapPointMatches, s, R, t, 7.5), src/LoopClosing.cc:388."""
    cal = default_cal(stereo=True)
    W = PhysWorld(seed, n_points, [pose(), pose(0.03, -0.01, t=(0.12, -0.02, 0.06))], cal, dup_frac=0.2, knife=[0, 1])
    S = _new_scene(SIM3, W, th, 0.75, True)
    S.fpar[2] = s12
    k1 = W.view(0, 1); k2 = W.view(1, 1)
    rng = W.rng
    _mps_for_view(S, W, k1, frac=0.85); _observe(S, k1, 0)
    _mps_for_view(S, W, k2, frac=0.85)
    for i in np.flatnonzero(k2.mp >= 0):
        S.mps[k2.mp[i]].obs.append((1, int(i)))
    for m in S.mps:
        m.bad = int(rng.random() < 0.03)
    # vpMatches12 from the earlier SearchByFboW / RANSAC inliers: KF1 keypoint -> a map point of KF2
    phys2 = {int(k2.phys[i]): int(k2.mp[i]) for i in np.flatnonzero(k2.mp >= 0)}
    m12 = np.full(k1.N, -1, np.int32)
    for i in np.flatnonzero(k1.mp >= 0):
        j = int(k1.phys[i])
        if j in phys2 and rng.random() < 0.25:
            m12[i] = phys2[j]
    S.views = [k1, k2]
    S.lists = [m12]
    T1, T2 = W.poses[0].astype(f64), W.poses[1].astype(f64)
    R12 = T1[:3, :3] @ T2[:3, :3].T
    t12 = -R12 @ T2[:3, 3] + T1[:3, 3]
    S.farr = [R12.astype(f32), (t12 * s12).astype(f32)]
    if other_cal:  # KF2 comes from another camera: the reference projects BOTH directions with pKF1's intrinsics
        S.fpar[4:8] = [cal["fx"] * 1.07, cal["fy"] * 0.95, cal["cx"] + 6.0, cal["cy"] - 4.0]
    return S


def _fuse_world(seed, n_points, n_views=4, stereo=True):
    """A keyframe (view 0) with its own map points and neighbours (views 1..) whose map points partly duplicate them."""
    cal = default_cal(stereo=stereo)
    poses = [pose(-0.03, t=(0.05, 0.02, -0.1))] + [pose(0.01 * k, t=(0.06 * k, -0.01 * k, 0.03 * k)) for k in range(1, n_views)]
    W = PhysWorld(seed, n_points, poses, cal, dup_frac=0.15, knife=[0, 1])
    views = [W.view(k, 1) for k in range(n_views)]
    return W, views


def scene_fuse(seed, th=3.0, n_points=1300, stereo=True):
    """LocalMapping::SearchInNeighbors -> Fuse(pKFi, vpMapPointMatches, 3.0), src/LocalMapping.cc:487 (and :512)."""
    W, views = _fuse_world(seed, n_points, stereo=stereo)
    S = _new_scene(FUSE, W, th, 0.6, True)
    rng = W.rng
    kf = views[0]
    # the keyframe's own points: observed in view 0 and (the same MapPoint object) in some neighbours
    own = {}
    for i in np.flatnonzero(kf.phys >= 0):
        if rng.random() < 0.55:
            j = int(kf.phys[i])
            kf.mp[i] = own[j] = len(S.mps)
            S.mps.append(W.map_point(j, W._flip(W.desc[j], rng.integers(0, 4))[0]))
    # the neighbours' points: a different MapPoint object for the same physical point (what Fuse merges), or the
    # keyframe's own object (IsInKeyFrame -> skipped)
    nb = {}
    for vid in range(1, len(views)):
        V = views[vid]
        for i in np.flatnonzero(V.phys >= 0):
            j = int(V.phys[i])
            r = rng.random()
            if j in own and r < 0.3:
                V.mp[i] = own[j]
            elif r < 0.8:
                if j not in nb:
                    nb[j] = len(S.mps)
                    S.mps.append(W.map_point(j, W._flip(W.desc[j], rng.integers(0, 4))[0]))
                V.mp[i] = nb[j]
    for vid, V in enumerate(views):
        seen = set()
        for i in np.flatnonzero(V.mp >= 0):          # one keypoint per map point per keyframe
            if int(V.mp[i]) in seen:
                V.mp[i] = -1
            else:
                seen.add(int(V.mp[i]))
        _observe(S, V, vid)
    for m in S.mps:
        m.bad = int(rng.random() < 0.03)
    S.views = views
    cand = sorted(set(nb.values()) | set(rng.choice(list(own.values()), 40)))
    S.lists = [rng.permutation(cand)]
    return S


def scene_fuse_sim3(seed, th=4.0, s=1.0, n_points=1300):
    """LoopClosing::SearchAndFuse -> Fuse(pKF, cvScw, mvpLoopMapPoints, 4, vpReplacePoints), src/LoopClosing.cc:633."""
    S = scene_fuse(seed, th=th, n_points=n_points)
    S.method = FUSE_SIM3
    T = S.views[0].Tcw
    Scw = np.eye(4, dtype=f32)
    Scw[:3, :3] = (f64(s) * T[:3, :3].astype(f64)).astype(f32); Scw[:3, 3] = (f64(s) * T[:3, 3].astype(f64)).astype(f32)
    S.farr = [Scw]
    return S


BOW_RELOC_BATCH = 12
METHOD_NAMES[BOW_RELOC_BATCH] = "SearchByFboW(candidate KFs, F) batched"


def scene_bow_reloc_batch(seed, K=5, nnratio=0.75, n_points=1000):
    """Tracking::Relocalization (src/Tracking.cc:1465-1490): ORBmatcher(0.75, true).SearchByFboW(pKF, mCurrentFrame, ...) for
    every candidate keyframe. The reference side loops over the single call; the drop-in side uses the batched overload."""
    blob, voc = _vocab()
    cal = default_cal(stereo=True)
    poses = [pose(0.03, t=(0.08, 0.0, 0.1), rz=0.05)] + [pose(0.01 * k, t=(0.05 * k, 0.0, 0.02 * k)) for k in range(K)]
    W = PhysWorld(seed, n_points, poses, cal, vocab_blob=blob, dup_frac=0.2)
    S = _new_scene(BOW_RELOC_BATCH, W, 0, nnratio, True)
    F = W.view(0, 0, roll=3.0)
    attach_featvec(F, voc)
    S.views = [F]
    for k in range(K):
        kf = W.view(k + 1, 1, detect=0.85 - 0.1 * k)   # ragged keyframe sizes
        _mps_for_view(S, W, kf, frac=0.8)
        _observe(S, kf, k + 1)
        attach_featvec(kf, voc)
        S.views.append(kf)
    for m in S.mps:
        m.bad = int(W.rng.random() < 0.04)
    return S


# ------------------------------------------------------------------------------------------------------------------
# Degenerate scenes: empty inputs of every search (the reference returns 0 matches and touches nothing).
# ------------------------------------------------------------------------------------------------------------------
def _empty_view(V):
    """A copy of view V with no keypoints at all."""
    E = View(np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8), np.zeros(0, f32), np.zeros(0, f32), V.Tcw, V.kind)
    return E


def _mutate(sc, fn):
    fn(sc)
    return sc


def _no_level0(sc):
    sc.views[0].keys["octave"][:] = 1
    sc.views[0].keys_un = sc.views[0].keys.copy()


def _all_have_points(sc):
    sc.views[0].mp[:] = np.arange(sc.views[0].N) % max(len(sc.mps), 1)


EDGE_SCENES = [
    ("frame_last_without_points", lambda: _mutate(scene_frame(0), lambda s: s.views[1].mp.fill(-1))),
    ("frame_current_without_keypoints", lambda: _mutate(scene_frame(1), lambda s: s.views.__setitem__(0, _empty_view(s.views[0])))),
    ("local_empty_list", lambda: _mutate(scene_local(0), lambda s: setattr(s, "lists", [np.zeros(0, np.int32)]))),
    ("local_nothing_in_view", lambda: _mutate(scene_local(1), lambda s: [setattr(m, "in_view", 0) for m in s.mps])),
    ("local_frame_without_keypoints", lambda: _mutate(scene_local(2), lambda s: s.views.__setitem__(0, _empty_view(s.views[0])))),
    ("reloc_kf_without_points", lambda: _mutate(scene_reloc(0), lambda s: s.views[1].mp.fill(-1))),
    ("kfproj_no_points", lambda: _mutate(scene_kfproj(0), lambda s: s.lists.__setitem__(0, np.zeros(0, np.int32)))),
    ("bow_kf_without_points", lambda: _mutate(scene_bow_kf_f(0), lambda s: s.views[0].mp.fill(-1))),
    ("bow_frame_empty_featvec", lambda: _mutate(scene_bow_kf_f(1), lambda s: setattr(s.views[1], "featvec", {}))),
    ("bow_kfkf_second_without_points", lambda: _mutate(scene_bow_kf_kf(0), lambda s: s.views[1].mp.fill(-1))),
    ("triang_all_keypoints_have_points", lambda: _mutate(scene_triang(0), _all_have_points)),
    ("sim3_kf2_without_points", lambda: _mutate(scene_sim3(0), lambda s: s.views[1].mp.fill(-1))),
    ("fuse_empty_list", lambda: _mutate(scene_fuse(0), lambda s: setattr(s, "lists", [np.zeros(0, np.int32)]))),
    ("fuse_sim3_empty_list", lambda: _mutate(scene_fuse_sim3(0), lambda s: setattr(s, "lists", [np.zeros(0, np.int32)]))),
    ("init_no_level0_keypoint", lambda: _mutate(scene_init(0), _no_level0)),
]
