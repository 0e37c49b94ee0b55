"""GPU parity of Frame::ComputeStereoMatches (device K9/K10) against the oracle's line-by-line restatement, both fed
the SAME keypoints/descriptors/pyramids (those produced by the GPU extractor, already proven identical to the oracle's)."""
import numpy as np
import pytest

from oracle_lib import OraclePrims, TUM1
from orbslam2_b200.synth import synth_stereo_pair

pytestmark = pytest.mark.gpu


# the last case has > 4096 left keypoints: the median-cut kernel then needs more than 48 KB of dynamic shared memory
@pytest.mark.parametrize("w,h,nf,bf,fx", [(1241, 376, 2000, 386.1448, 718.856), (752, 480, 1200, 47.9, 435.2),
                                          (1920, 1080, 6000, 386.1448, 1400.0)])
def test_stereo_vs_oracle(w, h, nf, bf, fx):
    import orbslam2_b200 as ob
    p = dict(TUM1); p["nfeatures"] = nf
    exL = ob.ORBextractor(**p); exR = ob.ORBextractor(**p)
    P = OraclePrims()
    pairs = [synth_stereo_pair(w, h, s) for s in (0, 1)]
    L = np.stack([a for a, _ in pairs]); R = np.stack([b for _, b in pairs])
    kL, dL, nL = exL.extract_batch(L); kR, dR, nR = exR.extract_batch(R)
    u, z = ob.compute_stereo_matches(exL, exR, bf, 0.0, fx)
    total = 0
    for f in range(2):
        pyrL = [exL.pyramid_level(l, f) for l in range(8)]; pyrR = [exR.pyramid_level(l, f) for l in range(8)]
        eu, ez, sad, n = P.stereo_match(kL[f, :nL[f]], dL[f, :nL[f]], kR[f, :nR[f]], dR[f, :nR[f]], pyrL, pyrR,
                                        exL.GetScaleFactors(), exL.GetInverseScaleFactors(), bf, 0.0, fx)
        assert np.array_equal(u[f, :nL[f]], eu), f"mvuRight frame {f}: {(u[f, :nL[f]] != eu).sum()} differ"
        assert np.array_equal(z[f, :nL[f]], ez), f"mvDepth frame {f}"
        total += int((eu >= 0).sum())
    assert total > 100, "synthetic pair should produce stereo matches"
    # single-frame entry point
    u1, z1 = ob.compute_stereo_matches(exL, exR, bf, 0.0, fx, frame=1)
    assert np.array_equal(u1[:nL[1]], u[1, :nL[1]]) and np.array_equal(z1[:nL[1]], z[1, :nL[1]])
