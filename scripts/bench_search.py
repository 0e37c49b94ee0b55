#!/usr/bin/env python
"""Latency of the device-resident ORBmatcher::SearchByProjection(Frame, Frame) (orbm_search_by_projection, host arrays
in and out, staging copies included) beside the oracle's sequential CPU restatement, on synthetic tracking pairs."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402


def main():
    import orbslam2_b200 as ob
    from orbslam2_b200.synth import synth_tracking_pair
    from oracle_lib import OraclePrims
    P = OraclePrims()
    m = ob.ORBmatcher(0.9, True)
    for n, th in ((1000, 15.0), (2000, 15.0), (8000, 15.0)):
        last, cur = synth_tracking_pair(n, 5, conflicts=n // 8)
        prm = m._proj_params(cur, last, th, False)
        for _ in range(3):
            got = m.SearchByProjection(cur, last, th, False)
        t0 = time.perf_counter()
        reps = 20
        for _ in range(reps):
            got = m.SearchByProjection(cur, last, th, False)
        gpu_ms = (time.perf_counter() - t0) / reps * 1e3
        t0 = time.perf_counter()
        for _ in range(5):
            exp = P.search_by_projection(prm, cur, last)
        cpu_ms = (time.perf_counter() - t0) / 5 * 1e3
        rounds, cands = m.search_stats()
        same = got[0] == exp[0] and all(np.array_equal(a, b) for a, b in zip(got[1:], exp[1:]))
        print(json.dumps({"metric": "search_by_projection_ms_per_frame_pair", "map_points": last.N, "keypoints": cur.N,
                          "th": th, "gpu_ms_host_call": gpu_ms, "cpu_oracle_ms_1_thread": cpu_ms, "nmatches": got[0],
                          "candidate_pairs": cands, "fixed_point_rounds": rounds, "identical_to_oracle": bool(same)}))

    # SearchByFboW(KF, F) and SearchForTriangulation on a synthetic vocabulary (k = 10, 5 levels), 2000 features per side
    from orbslam2_b200.synth import synth_vocabulary, synth_vocabulary_features
    from oracle_lib import oracle_search_by_bow
    blob = synth_vocabulary(k=10, levels=5, seed=31)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    rng = np.random.default_rng(2)
    for n in (1000, 2000, 8000):
        da = synth_vocabulary_features(blob, n, seed=5)
        pick = rng.permutation(n)[: int(0.7 * n)]
        db = np.concatenate([da[pick], synth_vocabulary_features(blob, n - len(pick), seed=6)])
        bits = np.unpackbits(db[:len(pick)], axis=1)
        for _ in range(6):
            bits[np.arange(len(pick)), rng.integers(0, 256, len(pick))] ^= 1
        db[:len(pick)] = np.packbits(bits, axis=1)
        db = np.ascontiguousarray(db[rng.permutation(n)])
        t0 = time.perf_counter()
        fa, fb = voc.transform(da, 4)[1], voc.transform(db, 4)[1]
        t_bow = (time.perf_counter() - t0) / 2 * 1e3
        A = ob.ORBmatcher.bow_side(da, (rng.random(n) * 360).astype(np.float32), fa, np.ones(n, np.uint8))
        B = ob.ORBmatcher.bow_side(db, (rng.random(n) * 360).astype(np.float32), fb, None)
        mb = ob.ORBmatcher(0.7, False)
        for _ in range(3):
            got = mb.SearchByFboW(A, B)
        t0 = time.perf_counter()
        for _ in range(20):
            got = mb.SearchByFboW(A, B)
        gpu_ms = (time.perf_counter() - t0) / 20 * 1e3
        t0 = time.perf_counter()
        for _ in range(5):
            exp = oracle_search_by_bow(A, B, 0.7, 50, False, False)
        cpu_ms = (time.perf_counter() - t0) / 5 * 1e3
        print(json.dumps({"metric": "search_by_bow_ms_per_keyframe_frame_pair", "features_per_side": n,
                          "gpu_ms_host_call": gpu_ms, "cpu_oracle_ms_1_thread": cpu_ms, "nmatches": got[0],
                          "fixed_point_rounds": mb.search_stats()[0], "transform_ms_per_frame_python_call": t_bow,
                          "identical_to_oracle": bool(got[0] == exp[0] and np.array_equal(got[1], exp[1]))}))
        mb.close()
    voc.close()

    # Relocalization shape: K candidate keyframes against ONE frame — K single calls vs one batched call
    # (orbm_search_by_bow_batch) vs the sequential CPU restatement
    voc = ob.Vocabulary(); voc.fromStream(blob)
    rng = np.random.default_rng(9)
    for n, K in ((1000, 20), (2000, 20)):
        dbF = synth_vocabulary_features(blob, n, seed=50)
        fbF = voc.transform(dbF, 4)[1]
        B = ob.ORBmatcher.bow_side(dbF, (rng.random(n) * 360).astype(np.float32), fbF, None)
        pairs = []
        for k in range(K):
            pick = rng.permutation(n)[: int(0.5 * n)]
            da = np.concatenate([dbF[pick], synth_vocabulary_features(blob, n - len(pick), seed=60 + k)])
            bits = np.unpackbits(da[:len(pick)], axis=1)
            for _ in range(6):
                bits[np.arange(len(pick)), rng.integers(0, 256, len(pick))] ^= 1
            da[:len(pick)] = np.packbits(bits, axis=1)
            da = np.ascontiguousarray(da[rng.permutation(n)])
            fa = voc.transform(da, 4)[1]
            pairs.append((ob.ORBmatcher.bow_side(da, (rng.random(n) * 360).astype(np.float32), fa, np.ones(n, np.uint8)), B))
        mb = ob.ORBmatcher(0.75, True)
        for _ in range(3):
            got = mb.SearchByFboWBatch(pairs)
        t0 = time.perf_counter()
        for _ in range(10):
            got = mb.SearchByFboWBatch(pairs)
        batch_ms = (time.perf_counter() - t0) / 10 * 1e3
        rounds = mb.search_stats()[0]
        t0 = time.perf_counter()
        for _ in range(5):
            single = [mb.SearchByFboW(a, b) for a, b in pairs]
        single_ms = (time.perf_counter() - t0) / 5 * 1e3
        t0 = time.perf_counter()
        exp = [oracle_search_by_bow(a, b, 0.75, 50, False, True) for a, b in pairs]
        cpu_ms = (time.perf_counter() - t0) * 1e3
        same = all(g[0] == e[0] and np.array_equal(g[1], e[1]) for g, e in zip(got, exp)) and \
            all(g[0] == e[0] and np.array_equal(g[1], e[1]) for g, e in zip(single, exp))
        print(json.dumps({"metric": "search_by_bow_ms_per_relocalization", "candidate_keyframes": K, "features_per_side": n,
                          "gpu_ms_batched_call": batch_ms, "gpu_ms_single_calls": single_ms, "cpu_oracle_ms_1_thread": cpu_ms,
                          "fixed_point_rounds": rounds, "nmatches_total": int(sum(g[0] for g in got)),
                          "identical_to_oracle": bool(same)}))
        mb.close()
    voc.close()

    # SearchForInitialization (monocular initialisation, window 100): host call vs the sequential CPU restatement
    import importlib.util
    spec = importlib.util.spec_from_file_location("tgs", os.path.join(ROOT, "tests", "test_gpu_search.py"))
    tgs = importlib.util.module_from_spec(spec); spec.loader.exec_module(tgs)
    from oracle_lib import oracle_search_for_initialization
    for n in (1000, 2000):
        F1, F2 = tgs._init_pair(3, n=n)
        mi = ob.ORBmatcher(0.9, True)
        start = np.ascontiguousarray(np.stack([F1.mvKeysUn["x"], F1.mvKeysUn["y"]], 1), np.float32)
        for _ in range(3):
            got = mi.SearchForInitialization(F1, F2, start.copy(), 100)
        t0 = time.perf_counter()
        for _ in range(20):
            prev = start.copy()
            got = mi.SearchForInitialization(F1, F2, prev, 100)
        gpu_ms = (time.perf_counter() - t0) / 20 * 1e3
        cands = mi.search_stats()[1]
        t0 = time.perf_counter()
        for _ in range(5):
            eprev = start.copy()
            exp = oracle_search_for_initialization(got[2], np.float32(0.9), F1, F2, eprev, 100)
        cpu_ms = (time.perf_counter() - t0) / 5 * 1e3
        print(json.dumps({"metric": "search_for_initialization_ms_per_frame_pair", "keypoints": n,
                          "level0_queries": int((F1.mvKeysUn["octave"] == 0).sum()), "candidate_pairs": cands,
                          "gpu_ms_host_call": gpu_ms, "cpu_oracle_ms_1_thread": cpu_ms, "nmatches": got[0],
                          "identical_to_oracle": bool(got[0] == exp[0] and np.array_equal(got[1], exp[1]) and
                                                      prev.tobytes() == eprev.tobytes())}))
        mi.close()

    # batched MapPoint::ComputeDistinctiveDescriptors: 20 000 map points with 3..40 observations
    rng = np.random.default_rng(1)
    counts = rng.integers(3, 41, 20000).astype(np.int32)
    off = np.zeros(len(counts) + 1, np.int32); off[1:] = np.cumsum(counts)
    desc = rng.integers(0, 256, (off[-1], 32), dtype=np.uint8)
    m.ComputeDistinctiveDescriptors(desc, off)
    t0 = time.perf_counter()
    for _ in range(5):
        bi, bm = m.ComputeDistinctiveDescriptors(desc, off)
    gpu_ms = (time.perf_counter() - t0) / 5 * 1e3
    t0 = time.perf_counter()
    ei, em = P.distinctive_descriptors(desc, off)
    cpu_ms = (time.perf_counter() - t0) * 1e3
    print(json.dumps({"metric": "distinctive_descriptors_ms_per_20000_map_points", "observations": int(off[-1]),
                      "gpu_ms_host_call": gpu_ms, "cpu_oracle_ms_1_thread": cpu_ms,
                      "identical_to_oracle": bool(np.array_equal(bi, ei) and np.array_equal(bm, em))}))


if __name__ == "__main__":
    main()
