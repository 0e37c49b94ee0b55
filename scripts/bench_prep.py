#!/usr/bin/env python
"""Cost of the upstream image prep fused into the ingest, 1 GPU, 384 frames per step:

  gray        the headline path (8-bit grey frames resident in HBM)
  bgr         BGR frames resident in HBM: cv::cvtColor(BGR2GRAY) + extraction
  rectified   752x480 grey frames + cv::remap(INTER_LINEAR) with EuRoC-like maps + extraction
  e2e_bgr     BGR frames in pinned host memory through ExtractorRing (3x the H2D bytes of the grey path)

and the CPU side of the same steps with cv2 (when importable: it is not on the GPU box) on one thread."""
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
    from orbslam2_b200.synth import synth_image
    from oracle_lib import TUM1, OraclePrep
    dev = torch.device("cuda", 0)
    B, steps = 384, 10

    def timed(ex, d_in):
        cap_w, cap_h = ex._out_size or (d_in.shape[2], d_in.shape[1])
        cap = ex.max_keypoints(cap_w, cap_h)
        out = (torch.empty((B, cap, 7), dtype=torch.float32, device=dev), torch.empty((B, cap, 32), dtype=torch.uint8, device=dev))
        stream = torch.cuda.ExternalStream(ex.cuda_stream(), device=dev)
        for _ in range(3):
            ex.extract_batch(d_in, out=out)
        torch.cuda.synchronize()
        l0 = ex.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            _, _, n = ex.extract_batch(d_in, out=out)
        e1.record(stream)
        torch.cuda.synchronize()
        return B * steps / (e0.elapsed_time(e1) * 1e-3), float(np.mean(n)), (ex.launch_count() - l0) // steps

    res = {}
    gray = np.stack([synth_image(640, 480, s) for s in range(8)])
    g = torch.from_numpy(gray[np.arange(B) % 8]).to(dev)
    ex = ob.ORBextractor(**TUM1)
    res["gray"] = timed(ex, g)
    ex.close()

    rng = np.random.default_rng(0)
    bgr = np.stack([np.clip(gray[i].astype(np.int32) * s // 8 + rng.integers(-6, 7, gray[i].shape), 0, 255)
                    for i in range(8) for s in (6, 8, 9)], 0).reshape(8, 3, 480, 640).transpose(0, 2, 3, 1).astype(np.uint8)
    bgr = np.ascontiguousarray(bgr)
    c = torch.from_numpy(bgr[np.arange(B) % 8]).to(dev)
    ex = ob.ORBextractor(**TUM1); ex.set_input_format("bgr")
    res["bgr"] = timed(ex, c)
    ex.close()

    src = np.stack([synth_image(752, 480, s) for s in range(8)])
    yy, xx = np.meshgrid(np.arange(480, dtype=np.float32), np.arange(752, dtype=np.float32), indexing="ij")
    r2 = ((xx - 376) ** 2 + (yy - 240) ** 2) / (376.0 ** 2)
    mx = ((xx - 376) * (1 + 0.05 * r2) + 376 + 1.3).astype(np.float32)
    my = ((yy - 240) * (1 + 0.05 * r2) + 240 - 0.8).astype(np.float32)
    s_ = torch.from_numpy(src[np.arange(B) % 8]).to(dev)
    ex = ob.ORBextractor(**{**TUM1, "nfeatures": 1200}); ex.set_rectify_maps(mx, my)
    res["rectified_752x480_1200kp"] = timed(ex, s_)
    ex.close()
    ex = ob.ORBextractor(**{**TUM1, "nfeatures": 1200})
    res["plain_752x480_1200kp"] = timed(ex, s_)
    ex.close()

    # end to end with BGR frames in pinned host memory
    host = torch.from_numpy(bgr[np.arange(B) % 8]).pin_memory()
    ring = ob.ExtractorRing(depth=2, chunk=96, **TUM1)
    for e in ring.ex:
        e.set_input_format("bgr")
    for _ in range(4):
        ring.push(host)
    ring.drain()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        ring.push(host)
    ring.drain()
    torch.cuda.synchronize()
    e2e_bgr = B * steps / (time.perf_counter() - t0)
    ring.close()

    cpu = None
    try:
        import cv2
        cv2.setNumThreads(1)
        t0 = time.perf_counter()
        for i in range(64):
            cv2.cvtColor(bgr[i % 8], cv2.COLOR_BGR2GRAY)
        t_cvt = (time.perf_counter() - t0) / 64
        t0 = time.perf_counter()
        for i in range(64):
            cv2.remap(src[i % 8], mx, my, cv2.INTER_LINEAR)
        t_remap = (time.perf_counter() - t0) / 64
        cpu = {"kind": "cv2 %s, 1 thread" % cv2.__version__, "cvtColor_ms_per_frame": t_cvt * 1e3, "remap_ms_per_frame": t_remap * 1e3}
    except ImportError:
        P = OraclePrep()
        t0 = time.perf_counter()
        for i in range(16):
            P.cvt_gray(bgr[i % 8], True)
        t_cvt = (time.perf_counter() - t0) / 16
        t0 = time.perf_counter()
        for i in range(16):
            P.remap(src[i % 8], mx, my)
        t_remap = (time.perf_counter() - t0) / 16
        cpu = {"kind": "oracle restatement (scalar C++), 1 thread", "cvtColor_ms_per_frame": t_cvt * 1e3, "remap_ms_per_frame": t_remap * 1e3}
    print(json.dumps({"metric": "orb_extract_frames_per_s_with_fused_prep", "frames_per_step": B,
                      "device": {k: {"frames_per_s": v[0], "keypoints_per_frame": v[1], "launches_per_step": v[2]} for k, v in res.items()},
                      "cvtcolor_added_us_per_frame": (1 / res["bgr"][0] - 1 / res["gray"][0]) * 1e6,
                      "remap_added_us_per_frame": (1 / res["rectified_752x480_1200kp"][0] - 1 / res["plain_752x480_1200kp"][0]) * 1e6,
                      "e2e_bgr_host_frames_per_s": e2e_bgr, "cpu_prep": cpu}))


if __name__ == "__main__":
    main()
