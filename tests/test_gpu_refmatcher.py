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
