#!/usr/bin/env python
"""Throughput of the device vocabulary-tree transform (orbv_transform_batch = Frame::ComputeFboW for a batch of frames)
on a synthetic ORB-sized vocabulary (k = 10, 6 levels, 10^6 words, level-4 feature vector), beside the UNMODIFIED
reference fbow::Vocabulary::transform (oracle/_ref/libref_fbow.so) on the host cores.

  device   descriptors resident in HBM (as orbx_extract_batch leaves them), results left in HBM; CUDA events on the
           vocabulary's stream
  host     descriptors and results in host memory (staging copies inside the call), wall clock
  cpu      the reference, one frame per call, on 1 thread and on all host threads (frames are independent)
"""
import argparse
import ctypes as C
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=384)
    ap.add_argument("--features", type=int, default=1000)
    ap.add_argument("--levels", type=int, default=6)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--cpu-frames", type=int, default=64)
    args = ap.parse_args()
    import torch
    import orbslam2_b200 as ob
    from orbslam2_b200 import _lib
    from orbslam2_b200.synth import synth_vocabulary, synth_vocabulary_features
    from oracle_lib import RefVocabulary, OracleVocabulary, have_ref_fbow

    t0 = time.perf_counter()
    blob = synth_vocabulary(k=10, levels=args.levels, seed=1)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    B, N = args.frames, args.features
    cap = N
    feats = synth_vocabulary_features(blob, B * N, seed=3).reshape(B, N, 32)
    n = np.full(B, N, np.int32)
    setup_s = time.perf_counter() - t0
    L = voc._L
    dev = torch.device("cuda", 0)
    d_desc = torch.from_numpy(feats).to(dev)
    d_u32 = [torch.empty((B, cap), dtype=torch.int32, device=dev) for _ in range(3)]
    d_ww = torch.empty((B, cap), dtype=torch.float32, device=dev)
    d_off = torch.empty((B, cap + 1), dtype=torch.int32, device=dev)
    nw, nn = np.zeros(B, np.int32), np.zeros(B, np.int32)
    stream = torch.cuda.ExternalStream(voc.cuda_stream(), device=dev)

    def step_device():
        voc._set_space(_lib.MEM_DEVICE, _lib.MEM_DEVICE)
        _lib.check(L.orbv_transform_batch(voc._h, d_desc.data_ptr(), n.ctypes.data, B, cap, 4, 0, d_u32[0].data_ptr(),
                                          d_ww.data_ptr(), nw.ctypes.data, d_u32[1].data_ptr(), d_off.data_ptr(),
                                          d_u32[2].data_ptr(), nn.ctypes.data))

    h = [np.zeros((B, cap), np.uint32) for _ in range(3)]
    h_ww, h_off = np.zeros((B, cap), np.float32), np.zeros((B, cap + 1), np.int32)

    def step_host():
        voc._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(L.orbv_transform_batch(voc._h, feats.ctypes.data, n.ctypes.data, B, cap, 4, 0, h[0].ctypes.data,
                                          h_ww.ctypes.data, nw.ctypes.data, h[1].ctypes.data, h_off.ctypes.data,
                                          h[2].ctypes.data, nn.ctypes.data))

    for _ in range(3):
        step_device()
    torch.cuda.synchronize()
    l0 = voc.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step_device()
    e1.record(stream)
    torch.cuda.synchronize()
    dev_s = e0.elapsed_time(e1) * 1e-3 / args.steps
    launches = voc.launch_count() - l0
    for _ in range(2):
        step_host()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    host_s = (time.perf_counter() - t0) / args.steps

    # parity of this very workload (frame 0 and the last frame) against the oracle
    from test_cpu_fbow import same
    orc = OracleVocabulary(blob)
    res = voc.transform_batch(feats[[0, B - 1]], n[:2], 4)
    identical = same(res[0], orc.transform(feats[0], 4)) and same(res[1], orc.transform(feats[B - 1], 4))

    cpu = None
    if have_ref_fbow():
        ref = RefVocabulary(blob)
        F = min(args.cpu_frames, B)
        t0 = time.perf_counter()
        for f in range(min(F, 8)):
            ref.transform(feats[f], 4)
        one = (time.perf_counter() - t0) / min(F, 8)
        threads = len(os.sched_getaffinity(0))
        refs = [RefVocabulary(blob) for _ in range(threads)]   # the reference object keeps per-call state: one each

        def work(t):
            for f in range(t, F * 4, threads):
                refs[t].transform(feats[f % B], 4)

        th = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        allc = (time.perf_counter() - t0) / (F * 4)
        cpu = {"kind": "reference", "frames_per_s_1_thread": 1.0 / one, "frames_per_s_all_threads": 1.0 / allc,
               "threads": threads, "sample": f"{F * 4} frames of {N} descriptors"}
    print(json.dumps({"metric": "fbow_transform_frames_per_s", "vocabulary": {"k": 10, "levels": args.levels,
                      "words": voc.getTotalWords(), "blocks": voc.size(), "bytes": len(blob)},
                      "frames_per_step": B, "features_per_frame": N, "store_level": 4,
                      "device_frames_per_s": B / dev_s, "device_features_per_s": B * N / dev_s,
                      "device_ms_per_step": dev_s * 1e3, "host_call_frames_per_s": B / host_s,
                      "host_call_ms_per_step": host_s * 1e3, "gpu_launches": launches,
                      "identical_to_oracle": bool(identical), "cpu_baseline": cpu, "setup_s": setup_s}))
    voc.close()


if __name__ == "__main__":
    main()
