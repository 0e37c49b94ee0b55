#!/usr/bin/env python
"""The per-frame front end of the Frame constructor chained on the device (src/Frame.cc:175-229: ExtractORB,
UndistortKeyPoints, ComputeFboW), 384 frames per step, nothing but the per-frame counts crossing PCIe:

    orbx_extract_batch (device in/out) -> orbx_undistort_keypoints (device out) -> orbv_transform_batch (device in/out)

Reports frames/s of the chain and of its parts (CUDA events on the extractor's stream; the three calls synchronise)."""
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
    from orbslam2_b200.synth import synth_image, synth_vocabulary
    from oracle_lib import TUM1
    dev = torch.device("cuda", 0)
    B, steps = 384, 10
    frames = np.stack([synth_image(640, 480, s) for s in range(8)])
    d_imgs = torch.from_numpy(frames[np.arange(B) % 8]).to(dev)
    ex = ob.ORBextractor(**TUM1)
    voc = ob.Vocabulary(); voc.fromStream(synth_vocabulary(k=10, levels=6, seed=1))
    cap = ex.max_keypoints(640, 480)
    kps = torch.empty((B, cap, 7), dtype=torch.float32, device=dev)
    kps_un = torch.empty_like(kps)
    desc = torch.empty((B, cap, 32), dtype=torch.uint8, device=dev)
    u32 = [torch.empty((B, cap), dtype=torch.int32, device=dev) for _ in range(3)]
    ww = torch.empty((B, cap), dtype=torch.float32, device=dev)
    off = torch.empty((B, cap + 1), dtype=torch.int32, device=dev)
    n, nw, nn = np.zeros(B, np.int32), np.zeros(B, np.int32), np.zeros(B, np.int32)
    K4 = np.array([517.306408, 516.469215, 318.643040, 255.313989], np.float32)      # TUM1.yaml
    D = np.array([0.262383, -0.953104, -0.005358, 0.002628, 1.163314], np.float32)
    L = ex._L
    ex._set_space(_lib.MEM_DEVICE, _lib.MEM_DEVICE)
    voc._set_space(_lib.MEM_DEVICE, _lib.MEM_DEVICE)
    t_parts = [0.0, 0.0, 0.0]

    def step(record=False):
        t0 = time.perf_counter()
        _lib.check(L.orbx_extract_batch(ex._h, d_imgs.data_ptr(), B, 640, 480, 640, 640 * 480, kps.data_ptr(), desc.data_ptr(),
                                        cap, n.ctypes.data))
        t1 = time.perf_counter()
        _lib.check(L.orbx_undistort_keypoints(ex._h, K4.ctypes.data, D.ctypes.data, len(D), kps_un.data_ptr(), cap))
        t2 = time.perf_counter()
        _lib.check(L.orbv_transform_batch(voc._h, desc.data_ptr(), n.ctypes.data, B, cap, 4, 0, u32[0].data_ptr(),
                                          ww.data_ptr(), nw.ctypes.data, u32[1].data_ptr(), off.data_ptr(),
                                          u32[2].data_ptr(), nn.ctypes.data))
        t3 = time.perf_counter()
        if record:
            for i, d in enumerate((t1 - t0, t2 - t1, t3 - t2)):
                t_parts[i] += d

    for _ in range(3):
        step()
    torch.cuda.synchronize()
    l0 = ex.launch_count() + voc.launch_count()
    t0 = time.perf_counter()
    for _ in range(steps):
        step(True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    print(json.dumps({"metric": "frame_front_end_frames_per_s", "chain": "extract -> undistort keypoints -> bag of words (level 4)",
                      "frames_per_step": B, "frames_per_s": B / dt, "ms_per_step": dt * 1e3,
                      "ms_extract": t_parts[0] / steps * 1e3, "ms_undistort": t_parts[1] / steps * 1e3,
                      "ms_bow": t_parts[2] / steps * 1e3, "keypoints_per_frame": float(n.mean()),
                      "words_per_frame": float(nw.mean()), "nodes_per_frame": float(nn.mean()),
                      "gpu_launches_per_step": (ex.launch_count() + voc.launch_count() - l0) // steps,
                      "host_bytes_per_step": int(3 * B * 4)}))
    ex.close(); voc.close()


if __name__ == "__main__":
    main()
