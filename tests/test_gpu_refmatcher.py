"""GPU parity against the reference's own code: every ORBmatcher search, as the C++ drop-in templates
(include/dropin/ORBmatcher.h over liborbx.so) execute it on the device, must reproduce what the UNMODIFIED
/root/reference/src/ORBmatcher.cc (oracle/_ref/scene_run_ref, prebuilt in the build container) returns on the same
scene: return value, output vectors and every MapPoint / KeyFrame mutation (tests/cpp/scene_run.cpp dumps them).
The scenes are deliberately ill-conditioned (PredictScale evaluated at integer +- ulp), so a last-bit difference in the
projection arithmetic shows up as different matches."""
import numpy as np
import pytest

import scene_lib as S
from test_cpu_refmatcher import SCENES, MIN_RET

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu_exe(tmp_path_factory):
    return S.build_gpu(tmp_path_factory.mktemp("scene_gpu"))


@pytest.mark.skipif(not S.have_ref_matcher(), reason="oracle/_ref/scene_run_ref not built")
@pytest.mark.parametrize("name,make", SCENES, ids=[n for n, _ in SCENES])
def test_dropin_on_gpu_vs_reference_method(gpu_exe, tmp_path, name, make):
    for seed in (0, 1, 2, 3):
        sc = make(seed)
        a = S.run_scene(S.REF_EXE, sc, tmp_path, "ref")
        b = S.run_scene(gpu_exe, sc, tmp_path, "gpu")
        pa = S.parse_dump(sc, a)
        assert pa["ret"] >= MIN_RET.get(name, 100)
        assert a.shape == b.shape and np.array_equal(a, b), \
            f"{S.METHOD_NAMES[sc.method]} seed {seed}: reference returned {pa['ret']}, GPU drop-in {int(b[0])}, " \
            f"{int((a != b).sum()) if a.shape == b.shape else 'shape'} dump words differ"


def test_keyframe_cache_reuses_staged_keyframe(tmp_path):
    """orbm_keyframe_hint: the second orbm_fuse_search against the same (immutable) keyframe takes its keypoints,
    descriptors and grid from the handle's cache — same results, one cache hit, four launches fewer; a different
    keyframe under another key misses."""
    import ctypes as C
    import orbslam2_b200 as ob
    from orbslam2_b200 import _lib
    from orbslam2_b200.synth import synth_tracking_pair
    kf, other = synth_tracking_pair(1200, 3, stereo=True)
    m = ob.ORBmatcher()
    L = m._L
    L.orbm_keyframe_hint.argtypes = [C.c_void_p, C.c_uint64]
    L.orbm_keyframe_cache_hits.restype = C.c_int64
    L.orbm_keyframe_cache_hits.argtypes = [C.c_void_p]
    rng = np.random.default_rng(1)
    n = 900
    src = rng.integers(0, kf.N, n)
    pts = dict(u=kf.mvKeysUn["x"][src] + rng.normal(0, 1, n).astype(np.float32), v=kf.mvKeysUn["y"][src] + rng.normal(0, 1, n).astype(np.float32),
               ur=np.zeros(n, np.float32), level=kf.mvKeysUn["octave"][src].astype(np.int32), valid=np.ones(n, np.uint8),
               desc=kf.mDescriptors[src].copy())
    base = m.FuseSearch(kf, pts, th=3.0, check_chi2=False)
    res, launches = [], []
    for key in (7, 7, 7, 9):
        _lib.check(L.orbm_keyframe_hint(m._h, key))
        l0 = m.launch_count()
        res.append(m.FuseSearch(kf if key == 7 else other, pts, th=3.0, check_chi2=False))
        launches.append(m.launch_count() - l0)
    assert L.orbm_keyframe_cache_hits(m._h) == 2
    # a miss builds the grid (one fused launch for a camera-sized keyframe), a hit does not
    assert launches[0] - launches[1] == 1 and launches[1] == launches[2] and launches[3] == launches[0]
    for r in res[:3]:
        assert np.array_equal(r[0], base[0]) and np.array_equal(r[1], base[1])
    assert not np.array_equal(res[3][0], base[0])
    m.close()


@pytest.mark.skipif(not S.have_ref_matcher(), reason="oracle/_ref/scene_run_ref not built")
@pytest.mark.parametrize("name,make", S.EDGE_SCENES, ids=[n for n, _ in S.EDGE_SCENES])
def test_degenerate_scenes_vs_reference(gpu_exe, tmp_path, name, make):
    """Empty inputs of every search (no map points, no keypoints, empty lists, empty feature vectors): the reference
    returns 0 and leaves the map untouched; so must the drop-in."""
    sc = make()
    a = S.run_scene(S.REF_EXE, sc, tmp_path, "ref")
    b = S.run_scene(gpu_exe, sc, tmp_path, "gpu")
    assert int(a[0]) == 0
    assert a.shape == b.shape and np.array_equal(a, b), f"{name}: reference returned {int(a[0])}, drop-in {int(b[0])}"
