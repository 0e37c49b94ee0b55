import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import orbslam2_b200 as ob
from oracle_lib import TUM1
from orbslam2_b200.synth import synth_image
ex = ob.ORBextractor(**TUM1)
k, d = ex(synth_image(640, 480, 0))
out = {f"c{l}": ex.candidates(l, 0) for l in range(8)}
out["k"] = k; out["d"] = d
np.savez(os.path.join(ROOT, "gpurun_out", "cands.npz"), **out)
print("dumped")
