"""The matcher / stereo half of the oracle pinned against the reference's own code (CPU, no GPU needed).

oracle/_ref/libref_orbmatcher.so is the UNMODIFIED /root/reference/src/ORBmatcher.cc plus the member bodies of Frame /
KeyFrame / MapPoint it calls (and Frame::ComputeStereoMatches, MapPoint::ComputeDistinctiveDescriptors), compiled where
they lie against oracle/shim_matcher (oracle/Makefile). These tests assert, per method,
    reference method on a scene  ==  drop-in template host half + the oracle's orc_* search on the same scene
(return value, every output vector, every MapPoint / KeyFrame mutation), and that the cv::Mat stand-in the reference is
compiled against computes what the real OpenCV computes (cv2 golden vectors). The GPU counterpart
(tests/test_gpu_refmatcher.py) runs the same scenes through liborbx.so.
"""
import ctypes as C
import os

import numpy as np
import pytest

import scene_lib as S
from oracle_lib import OraclePrims, Oracle, canonical, TUM1, _p

needs_ref = pytest.mark.skipif(not S.have_ref_matcher(), reason="oracle/_ref/libref_orbmatcher.so not built (no /root/reference)")
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cv2_matops.npz")


@pytest.fixture(scope="module")
def ref():
    L = C.CDLL(S.REF_SO)
    L.refm_mat_eval.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.refm_descriptor_distance.argtypes = [C.c_void_p, C.c_void_p]
    L.refm_three_maxima.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    L.refm_distinctive_descriptors.argtypes = [C.c_void_p] * 2 + [C.c_int] + [C.c_void_p] * 2
    L.refm_stereo_match.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p] + [C.c_void_p] * 6 + \
                                   [C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
    return L


@pytest.fixture(scope="module")
def hostcheck(tmp_path_factory):
    return S.build_hostcheck(tmp_path_factory.mktemp("scene_host"))


def _eval(L, op, A, B=None, Cm=None):
    A = np.ascontiguousarray(A, np.float32)
    out = np.zeros(16, np.float32); sc = C.c_double()
    B = None if B is None else np.ascontiguousarray(B, np.float32)
    Cm = None if Cm is None else np.ascontiguousarray(Cm, np.float32)
    n = L.refm_mat_eval(op, _p(A), A.shape[0], A.shape[1], None if B is None else _p(B), 0 if B is None else B.shape[0],
                        0 if B is None else B.shape[1], None if Cm is None else _p(Cm), _p(out), C.byref(sc))
    return out[:n].copy(), sc.value


@needs_ref
def test_cv_mat_shim_matches_opencv(ref):
    """Every product form the reference uses, bit for bit against cv2.gemm / cv2.norm (OpenCV 4.13)."""
    g = np.load(GOLD)
    assert str(g["cv2_version"]).startswith("4.")
    for name in ("33_31", "33_33", "44_44", "44_41", "13_33", "34_41"):
        A, B, Cm = g[f"mul_{name}_A"], g[f"mul_{name}_B"], g[f"mul_{name}_C"]
        for i in range(len(A)):
            assert _eval(ref, 0, A[i], B[i])[0].tobytes() == g[f"mul_{name}_D"][i].tobytes(), (name, i)
            assert _eval(ref, 1, A[i], B[i], Cm[i])[0].tobytes() == g[f"mul_{name}_DC"][i].tobytes(), (name, i)
            assert _eval(ref, 4, A[i], B[i])[0].tobytes() == g[f"mul_{name}_DN"][i].tobytes(), (name, i)
    for name in ("33_31", "33_33"):
        A, B = g[f"tmul_{name}_A"], g[f"tmul_{name}_B"]
        for i in range(len(A)):
            assert _eval(ref, 2, A[i], B[i])[0].tobytes() == g[f"tmul_{name}_D"][i].tobytes(), (name, i)
            assert _eval(ref, 3, A[i], B[i])[0].tobytes() == g[f"tmul_{name}_DN"][i].tobytes(), (name, i)
    for i in range(len(g["norm_A"])):
        assert _eval(ref, 5, g["norm_A"][i])[1] == g["norm_L2"][i]
        assert _eval(ref, 6, g["l1_A"][i], g["l1_B"][i])[1] == g["l1"][i]
    # the small-matrix path really is float accumulation: the double-accumulated product differs on most inputs
    A, B = g["mul_33_31_A"], g["mul_33_31_B"]
    dbl = np.stack([(A[i].astype(np.float64) @ B[i].astype(np.float64)).astype(np.float32) for i in range(len(A))])
    assert (dbl != g["mul_33_31_D"]).any(axis=(1, 2)).mean() > 0.3


@needs_ref
def test_descriptor_distance_and_three_maxima(ref):
    rng = np.random.default_rng(3)
    P = OraclePrims()
    d = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    d[0] = 0; d[1] = 255
    for i in range(0, 299):
        a, b = np.ascontiguousarray(d[i]), np.ascontiguousarray(d[i + 1])
        assert ref.refm_descriptor_distance(_p(a), _p(b)) == P.hamming(a, b) == int(np.unpackbits(a ^ b).sum())
    # ComputeThreeMaxima against a direct restatement of its rule (src/ORBmatcher.cc:1597-1638)
    for t in range(300):
        cnt = rng.integers(0, [3, 12, 60][t % 3], 30).astype(np.int32)
        out = np.zeros(3, np.int32)
        ref.refm_three_maxima(_p(cnt), 30, _p(out))
        m = [0, 0, 0]; ind = [-1, -1, -1]
        for i, s in enumerate(cnt):
            if s > m[0]:
                m, ind = [s, m[0], m[1]], [i, ind[0], ind[1]]
            elif s > m[1]:
                m, ind = [m[0], s, m[1]], [ind[0], i, ind[1]]
            elif s > m[2]:
                m[2], ind[2] = s, i
        if m[1] < 0.1 * np.float32(m[0]):
            ind[1] = ind[2] = -1
        elif m[2] < 0.1 * np.float32(m[0]):
            ind[2] = -1
        assert list(out) == ind, (cnt, out, ind)


@needs_ref
def test_distinctive_descriptors_oracle_vs_reference(ref):
    """orc_distinctive_descriptors == MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307), ragged
    observation counts 0..40 with duplicated rows (ties on the median)."""
    rng = np.random.default_rng(8)
    P = OraclePrims()
    cnt = rng.integers(0, 41, 400); cnt[:4] = [0, 1, 2, 3]
    off = np.zeros(401, np.int32); off[1:] = np.cumsum(cnt)
    base = rng.integers(0, 256, (400, 32), dtype=np.uint8)
    rows = np.repeat(base, cnt, axis=0)
    bits = np.unpackbits(rows, axis=1)
    flips = rng.random(bits.shape) < 0.06
    flips[rng.random(len(bits)) < 0.3] = False       # exact duplicates
    desc = np.packbits(bits ^ flips, axis=1)
    bi, _ = P.distinctive_descriptors(desc, off)
    out = np.zeros((400, 32), np.uint8); chosen = np.zeros(400, np.int32)
    ref.refm_distinctive_descriptors(_p(desc), _p(off), 400, _p(out), _p(chosen))
    for p in range(400):
        if cnt[p] == 0:
            assert bi[p] == -1 and chosen[p] == 0
        else:
            assert chosen[p] == 1 and np.array_equal(out[p], desc[off[p] + bi[p]]), p


@needs_ref
@pytest.mark.parametrize("w,h,nf,seed", [(640, 480, 1000, 5), (752, 480, 1200, 6), (1241, 376, 2000, 7)])
def test_stereo_oracle_vs_reference(ref, w, h, nf, seed):
    """orc_stereo_match == Frame::ComputeStereoMatches (src/Frame.cc:464-642) on extracted stereo pairs: mvuRight and
    mvDepth bit-identical."""
    from orbslam2_b200.synth import synth_stereo_pair
    left, right = synth_stereo_pair(w, h, seed)
    par = dict(TUM1); par["nfeatures"] = nf
    oL, oR = Oracle(**par), Oracle(**par)
    kL, dL = canonical(*oL.extract(left, 4 * nf)); kR, dR = canonical(*oR.extract(right, 4 * nf))
    pyrL = [oL.level(l) for l in range(8)]; pyrR = [oR.level(l) for l in range(8)]
    t, _, _ = oL.tables()
    mbf, mb = np.float32(40.0), np.float32(40.0 / 520.0)
    P = OraclePrims()
    eu, ez, _, nacc = P.stereo_match(kL, dL, kR, dR, pyrL, pyrR, t[0], t[1], mbf, 0.0, float(mbf / mb))
    nl = 8
    PL = (C.c_void_p * nl)(*[p.ctypes.data for p in pyrL]); PR = (C.c_void_p * nl)(*[p.ctypes.data for p in pyrR])
    lw = np.array([p.shape[1] for p in pyrL], np.int32); lh = np.array([p.shape[0] for p in pyrL], np.int32)
    pl = np.array([p.strides[0] for p in pyrL], np.int32); pr = np.array([p.strides[0] for p in pyrR], np.int32)
    sc = np.ascontiguousarray(t[0], np.float32); isc = np.ascontiguousarray(t[1], np.float32)
    ur = np.zeros(len(kL), np.float32); dp = np.zeros(len(kL), np.float32)
    kLc, kRc = np.ascontiguousarray(kL), np.ascontiguousarray(kR)
    ref.refm_stereo_match(_p(kLc), len(kL), _p(dL), _p(kRc), len(kR), _p(dR), PL, PR, _p(lw), _p(lh), _p(pl), _p(pr), nl,
                          _p(sc), _p(isc), mbf, mb, _p(ur), _p(dp))
    assert ur.tobytes() == eu.tobytes() and dp.tobytes() == ez.tobytes()
    assert (ur >= 0).sum() > 100


SCENES = [
    ("frame_stereo", lambda s: S.scene_frame(s)),
    ("frame_mono", lambda s: S.scene_frame(s + 10, stereo=False, bMono=True, th=7.0)),
    ("frame_noori", lambda s: S.scene_frame(s + 20, th=7.0, check_ori=False, motion=(0.02, 0.0, -0.15))),
    ("local", lambda s: S.scene_local(s)),
    ("local_mono_th5", lambda s: S.scene_local(s + 20, stereo=False, th=5.0, nnratio=0.9)),
    ("local_th1", lambda s: S.scene_local(s + 30, th=1.0)),
    ("reloc", lambda s: S.scene_reloc(s)),
    ("reloc_th3", lambda s: S.scene_reloc(s + 10, th=3.0, orb_dist=64)),
    ("kfproj", lambda s: S.scene_kfproj(s)),
    ("kfproj_s1.1", lambda s: S.scene_kfproj(s + 10, th=4, s=1.1, dup=0.6)),
    ("init", lambda s: S.scene_init(s)),
    ("init_w30", lambda s: S.scene_init(s + 10, window=30, nnratio=0.7, check_ori=False)),
    ("bow_kf_f", lambda s: S.scene_bow_kf_f(s)),
    ("bow_kf_f_noori", lambda s: S.scene_bow_kf_f(s + 10, nnratio=0.75, check_ori=False)),
    ("bow_kf_kf", lambda s: S.scene_bow_kf_kf(s)),
    ("triang", lambda s: S.scene_triang(s)),
    ("triang_onlystereo", lambda s: S.scene_triang(s + 10, only_stereo=True)),
    ("triang_mono_ori", lambda s: S.scene_triang(s + 20, stereo=False, check_ori=True)),
    ("sim3", lambda s: S.scene_sim3(s)),
    ("sim3_s1.05", lambda s: S.scene_sim3(s + 10, s12=1.05, th=5.0)),
    ("sim3_two_cameras", lambda s: S.scene_sim3(s + 20, other_cal=True)),
    ("fuse", lambda s: S.scene_fuse(s)),
    ("fuse_mono", lambda s: S.scene_fuse(s + 10, stereo=False, th=5.0)),
    ("fuse_sim3", lambda s: S.scene_fuse_sim3(s)),
    ("fuse_sim3_s.9", lambda s: S.scene_fuse_sim3(s + 10, s=0.9, th=6.0)),
    ("bow_reloc_batch", lambda s: S.scene_bow_reloc_batch(s)),
]
MIN_RET = {"sim3_s1.05": 50, "sim3_two_cameras": 20}


@needs_ref
@pytest.mark.parametrize("name,make", SCENES, ids=[n for n, _ in SCENES])
def test_reference_method_vs_hostcheck(hostcheck, tmp_path, name, make):
    """scene_run_ref (unmodified src/ORBmatcher.cc) == drop-in host half + oracle search, whole dump identical."""
    for seed in (0, 1, 2):
        sc = make(seed)
        a = S.run_scene(S.REF_EXE, sc, tmp_path, "ref")
        b = S.run_scene(hostcheck, sc, tmp_path, "host")
        pa = S.parse_dump(sc, a)
        assert pa["ret"] >= MIN_RET.get(name, 100), f"{name}: scene too easy ({pa['ret']} matches)"
        assert a.shape == b.shape and np.array_equal(a, b), \
            f"{S.METHOD_NAMES[sc.method]} seed {seed}: reference returned {pa['ret']}, host half + oracle {int(b[0])}"


@pytest.mark.skipif(not S.have_ref_matcher(), reason="oracle/_ref/scene_run_ref not built")
@pytest.mark.parametrize("name,make", S.EDGE_SCENES, ids=[n for n, _ in S.EDGE_SCENES])
def test_degenerate_scenes_vs_reference(hostcheck, tmp_path, name, make):
    """Empty inputs of every search (no map points, no keypoints, empty lists, empty feature vectors): the reference
    returns 0 and leaves the map untouched; so must the drop-in."""
    sc = make()
    a = S.run_scene(S.REF_EXE, sc, tmp_path, "ref")
    b = S.run_scene(hostcheck, sc, tmp_path, "host")
    assert int(a[0]) == 0
    assert a.shape == b.shape and np.array_equal(a, b), f"{name}: reference returned {int(a[0])}, drop-in {int(b[0])}"
