import os, sys, importlib.util
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import orbslam2_b200 as ob
spec = importlib.util.spec_from_file_location("tgs", os.path.join(ROOT, "tests", "test_gpu_search.py"))
tgs = importlib.util.module_from_spec(spec); spec.loader.exec_module(tgs)
for n in (1000, 2000):
    F1, F2 = tgs._init_pair(3, n=n)
    mi = ob.ORBmatcher(0.9, True)
    start = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
    for _ in range(3):
        mi.SearchForInitialization(F1, F2, start.copy(), 100)
    mi.close()
