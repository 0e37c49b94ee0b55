#!/usr/bin/env python
"""Single-frame latency of the drop-in call (one 640x480 frame per orbx_extract call, host buffers in and out, the way
Frame::ExtractORB uses the extractor), beside the reference's CPU extractor on one thread."""
import ctypes as C
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402


def main():
    import torch
    import orbslam2_b200 as ob
    from orbslam2_b200 import _lib
    from orbslam2_b200.synth import synth_image
    from oracle_lib import TUM1, RefExtractor, have_ref
    res = {}
    for (w, h, nf) in ((640, 480, 1000), (752, 480, 1200), (1241, 376, 2000)):
        p = dict(TUM1); p["nfeatures"] = nf
        ex = ob.ORBextractor(**p)
        imgs = [torch.from_numpy(synth_image(w, h, s)).pin_memory() for s in range(4)]
        cap = ex.max_keypoints(w, h)
        kps = torch.empty((cap, 7), dtype=torch.float32).pin_memory()
        desc = torch.empty((cap, 32), dtype=torch.uint8).pin_memory()
        n = C.c_int()
        L = ex._L

        def one(i):
            _lib.check(L.orbx_extract(ex._h, imgs[i % 4].data_ptr(), w, h, w, kps.data_ptr(), desc.data_ptr(), cap, C.byref(n)))

        for i in range(20):
            one(i)
        l0 = ex.launch_count()
        t = []
        for i in range(200):
            t0 = time.perf_counter()
            one(i)
            t.append(time.perf_counter() - t0)
        t = np.array(t) * 1e3
        r = {"gpu_ms_median": float(np.median(t)), "gpu_ms_p95": float(np.percentile(t, 95)), "keypoints": n.value,
             "launches_per_frame": (ex.launch_count() - l0) / 200}
        if have_ref():
            ref = RefExtractor(**p)
            img = synth_image(w, h, 0)
            ref.extract(img, 4 * nf)
            t0 = time.perf_counter()
            for _ in range(5):
                ref.extract(img, 4 * nf)
            r["cpu_reference_ms_1_thread"] = (time.perf_counter() - t0) / 5 * 1e3
        res[f"{w}x{h}_{nf}kp"] = r
        ex.close()
    print(json.dumps({"metric": "orb_extract_single_frame_latency_ms", "host_buffers": "pinned", "results": res}))


if __name__ == "__main__":
    main()
