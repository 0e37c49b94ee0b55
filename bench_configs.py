"""bench_configs.py — `bench.py --config {2,3,4,5}`: the other BASELINE.json configs as driver-runnable lines with the same
schema as the headline (value, e2e with host buffers, roofline, cpu_baseline, parity, clocks, gpu_launches).

  2  KITTI-shape stereo 1241x376, 2000 kp: extraction of both images + ComputeStereoMatches          pairs/s, weak
  3  EuRoC-shape stereo 752x480, 1200 kp: 256 pairs sharded over the ranks (sharding.shard_range)    pairs/s, STRONG
  4  brute-force Hamming 2-NN (k = 2, ratio 0.7) 100k x 1M, queries sharded over the ranks           Gpairs/s, strong
  5  4K 3840x2160 frames, 8000 kp                                                                   frames/s, weak

Every run starts with a PARITY GATE: before anything is timed, one frame / stereo pair of the workload
(config 4: 512 sampled queries of the full 100k x 1M problem) goes through the CUDA path and through the CPU oracle and
the outputs are compared bit for bit; the verdict is the line's "parity" object and a failed gate makes the run exit 1
after printing the line. `--impl reference` times the reference's CPU code for the same config on the host cores.
"""
import json
import os
import sys
import threading
import time

import numpy as np

import bench as B

CFG = {
    2: dict(name="kitti_stereo_1241x376_2000kp", w=1241, h=376, nf=2000, stereo=True, bf=386.1448, fx=718.856,
            per_gpu=64, total=None, distinct=4, metric="orb_stereo_pairs_per_s_1241x376_2000kp", unit="pairs/s"),
    3: dict(name="euroc_stereo_752x480_1200kp_256_pairs_sharded", w=752, h=480, nf=1200, stereo=True, bf=47.90639384423901,
            fx=435.2046959714599, per_gpu=None, total=256, distinct=4, metric="orb_stereo_pairs_per_s_752x480_1200kp",
            unit="pairs/s"),
    5: dict(name="uhd_3840x2160_8000kp", w=3840, h=2160, nf=8000, stereo=False, bf=0.0, fx=1.0, per_gpu=16, total=None,
            distinct=2, metric="orb_extract_frames_per_s_3840x2160_8000kp", unit="frames/s"),
}


def level_area_sum(w, h, nlevels=8, scale=1.2):
    total, sf = 0, np.float32(1.0)
    for _ in range(nlevels):
        inv = np.float32(1.0) / sf
        total += int(np.rint(np.float32(w) * inv)) * int(np.rint(np.float32(h) * inv))
        sf = np.float32(np.float64(sf) * np.float64(np.float32(scale)))
    return total


def params_for(cfg):
    p = dict(B.PARAMS)
    p["nfeatures"] = cfg["nf"]
    return p


def synth_inputs(cfg, seed0, count):
    """`count` frames (mono) or (left, right) pairs, tiled from cfg['distinct'] distinct synthetic ones."""
    from orbslam2_b200.synth import synth_image, synth_stereo_pair
    w, h, d = cfg["w"], cfg["h"], cfg["distinct"]
    if cfg["stereo"]:
        ps = [synth_stereo_pair(w, h, seed0 + s) for s in range(d)]
        idx = np.arange(count) % d
        return (np.ascontiguousarray(np.stack([ps[i][0] for i in idx])), np.ascontiguousarray(np.stack([ps[i][1] for i in idx])))
    fr = [synth_image(w, h, seed0 + s) for s in range(d)]
    return (np.ascontiguousarray(np.stack([fr[i % d] for i in range(count)])), None)


# ------------------------------------------------------------------------------------------------ CPU side
class CpuPairPool:
    """The reference's CPU path for one config on every host core: ORBextractor (oracle/_ref/libref_orbextractor.so, the
    unmodified src/ORBextractor.cc, else the oracle port) on the left (and right) image, then Frame::ComputeStereoMatches
    (oracle/_ref/libref_orbmatcher.so, the unmodified src/Frame.cc:464-642, else the oracle port)."""

    def __init__(self, cfg, threads=None):
        import ctypes as C
        import oracle_lib
        self.cfg = cfg
        self.threads = threads or B.cpu_cores()
        have_ref = os.path.exists(oracle_lib.REF_SO)
        cls = oracle_lib.RefExtractor if have_ref else oracle_lib.Oracle
        self.kind = "reference" if have_ref else "port"
        p = params_for(cfg)
        self.exL = [cls(**p) for _ in range(self.threads)]
        self.exR = [cls(**p) for _ in range(self.threads)] if cfg["stereo"] else None
        self.prims = oracle_lib.OraclePrims()
        self.refm = None
        so = os.path.join(B.ROOT, "oracle", "_ref", "libref_orbmatcher.so")
        if cfg["stereo"] and os.path.exists(so):
            self.refm = C.CDLL(so)
            self.refm.refm_stereo_match.argtypes = ([C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p] +
                                                    [C.c_void_p] * 6 + [C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float,
                                                                        C.c_void_p, C.c_void_p])
        if cfg["stereo"] and self.refm is None:
            self.kind = "port"

    def one(self, t, left, right):
        """Returns (kL, dL, uRight, depth) of one frame / pair on thread slot t."""
        import ctypes as C
        from oracle_lib import canonical, _p
        cfg = self.cfg
        cap = 4 * cfg["nf"]
        kL, dL = self.exL[t].extract(left, cap)
        if not cfg["stereo"]:
            return kL, dL, None, None
        kR, dR = self.exR[t].extract(right, cap)
        pyrL = [self.exL[t].level(l) for l in range(8)]; pyrR = [self.exR[t].level(l) for l in range(8)]
        tab, _, _ = self.exL[t].tables()
        mbf, mb = np.float32(cfg["bf"]), np.float32(np.float32(cfg["bf"]) / np.float32(cfg["fx"]))
        if self.refm is None:
            u, z, _, _ = self.prims.stereo_match(kL, dL, kR, dR, pyrL, pyrR, tab[0], tab[1], mbf, 0.0, float(np.float32(mbf / mb)))
            return kL, dL, u, z
        nl = 8
        PL = (C.c_void_p * nl)(*[q.ctypes.data for q in pyrL]); PR = (C.c_void_p * nl)(*[q.ctypes.data for q in pyrR])
        lw = np.array([q.shape[1] for q in pyrL], np.int32); lh = np.array([q.shape[0] for q in pyrL], np.int32)
        pl = np.array([q.strides[0] for q in pyrL], np.int32); pr = np.array([q.strides[0] for q in pyrR], np.int32)
        sc = np.ascontiguousarray(tab[0], np.float32); isc = np.ascontiguousarray(tab[1], np.float32)
        u = np.zeros(len(kL), np.float32); z = np.zeros(len(kL), np.float32)
        kLc, kRc = np.ascontiguousarray(kL), np.ascontiguousarray(kR)
        dLc, dRc = np.ascontiguousarray(dL), np.ascontiguousarray(dR)
        self.refm.refm_stereo_match(_p(kLc), len(kL), _p(dLc), _p(kRc), len(kR), _p(dRc), PL, PR, _p(lw), _p(lh), _p(pl),
                                    _p(pr), nl, _p(sc), _p(isc), mbf, mb, _p(u), _p(z))
        return kL, dL, u, z

    def run(self, left, right, count):
        def work(t):
            for i in range(t, count, self.threads):
                self.one(t, left[i % len(left)], None if right is None else right[i % len(right)])

        th = [threading.Thread(target=work, args=(t,)) for t in range(self.threads)]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        return time.perf_counter() - t0


def cpu_baseline(cfg, left, right, seconds):
    pool = CpuPairPool(cfg)
    dt = pool.run(left, right, pool.threads)
    rounds = int(max(1, min(32, (seconds - dt) / max(dt, 1e-3))))
    count = pool.threads * rounds
    dt = pool.run(left, right, count)
    what = "stereo pairs (2 extractions + ComputeStereoMatches)" if cfg["stereo"] else "frames"
    return {"value": count / dt, "unit": cfg["unit"], "cores": pool.threads, "kind": pool.kind,
            "sample": f"{count} synthetic {cfg['w']}x{cfg['h']} {what}, {pool.threads} threads x {rounds} each, {dt:.1f} s wall"}


# ------------------------------------------------------------------------------------------------ parity gates
def parity_shape(cfg, ob, local, left, right):
    """One frame / pair through the CUDA path (host buffers, the public call) and through the CPU oracle."""
    from oracle_lib import Oracle, OraclePrims, canonical
    p = params_for(cfg)
    w, h = cfg["w"], cfg["h"]
    exL = ob.ORBextractor(**p, device=local)
    kG, dG, nG = exL.extract_batch(left[:1])
    kG, dG = kG[0, :nG[0]], dG[0, :nG[0]]
    oL = Oracle(**p)
    kO, dO = canonical(*oL.extract(left[0], exL.max_keypoints(w, h)))
    res = {"what": f"frame 0 of the workload ({w}x{h}, {cfg['nf']} kp) vs the CPU oracle (oracle/liborb_oracle.so)",
           "keypoints_gpu": int(nG[0]), "keypoints_oracle": int(len(kO))}
    same_n = len(kO) == nG[0]
    res["keypoint_records_identical"] = bool(same_n and kG.tobytes() == kO.tobytes())
    if same_n and not res["keypoint_records_identical"]:
        res["max_abs_xy_diff"] = float(max(np.abs(kG["x"] - kO["x"]).max(), np.abs(kG["y"] - kO["y"]).max()))
        res["max_abs_angle_diff_deg"] = float(np.abs(kG["angle"] - kO["angle"]).max())
    res["descriptors_identical_frac"] = float((dG == dO).all(axis=1).mean()) if same_n and len(kO) else 0.0
    ok = same_n and res["descriptors_identical_frac"] >= 0.9999 and (
        res["keypoint_records_identical"] or (res["max_abs_xy_diff"] <= 1e-4 and res["max_abs_angle_diff_deg"] <= 1e-2))
    if cfg["stereo"]:
        exR = ob.ORBextractor(**p, device=local)
        exL.extract_batch(left[:1]); exR.extract_batch(right[:1])
        mbf = np.float32(cfg["bf"]); mb = np.float32(mbf / np.float32(cfg["fx"]))
        u, z = ob.compute_stereo_matches(exL, exR, float(mbf), 0.0, float(np.float32(mbf / mb)), frame=0)
        oR = Oracle(**p)
        kR, dR = canonical(*oR.extract(right[0], exL.max_keypoints(w, h)))
        t, _, _ = oL.tables()
        eu, ez, _, _ = OraclePrims().stereo_match(kO, dO, kR, dR, [oL.level(l) for l in range(8)], [oR.level(l) for l in range(8)],
                                                  t[0], t[1], mbf, 0.0, float(np.float32(mbf / mb)))
        res["stereo_uright_depth_identical"] = bool(same_n and np.array_equal(u[:len(eu)], eu) and np.array_equal(z[:len(ez)], ez))
        res["stereo_matches"] = int((eu >= 0).sum())
        ok = ok and res["stereo_uright_depth_identical"]
        exR.close()
    exL.close()
    res["ok"] = bool(ok)
    return res


# ------------------------------------------------------------------------------------------------ GPU side, configs 2/3/5
def run_shape(args, cfg, strong):
    out = B.StdoutGuard()
    import torch
    import torch.distributed as dist
    import orbslam2_b200 as ob
    from orbslam2_b200 import sharding

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = B.bind_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev, timeout=__import__("datetime").timedelta(seconds=300))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=op)
        return float(t.item())

    vmax = lambda x: reduce(x, dist.ReduceOp.MAX)  # noqa: E731
    vsum = lambda x: reduce(x, dist.ReduceOp.SUM)  # noqa: E731

    w, h, nf, stereo = cfg["w"], cfg["h"], cfg["nf"], cfg["stereo"]
    if strong:
        lo, hi = sharding.shard_range(cfg["total"], rank, world)   # contiguous block of the 256 pairs
        n_local, n_total = hi - lo, cfg["total"]
    else:
        n_local, n_total = cfg["per_gpu"], cfg["per_gpu"] * world
    left, right = synth_inputs(cfg, rank * cfg["distinct"], n_local)
    parity = parity_shape(cfg, ob, local, left, right) if rank == 0 else None

    p = params_for(cfg)
    exL = ob.ORBextractor(**p, device=local)
    exR = ob.ORBextractor(**p, device=local) if stereo else None
    cap = exL.max_keypoints(w, h)
    hL = torch.from_numpy(left).pin_memory()
    hR = torch.from_numpy(right).pin_memory() if stereo else None
    dL = hL.to(dev); dR = hR.to(dev) if stereo else None

    def outs(device):
        kw = dict(device=device) if device is not None else {}
        k = torch.empty((n_local, cap, 7), dtype=torch.float32, **kw)
        d = torch.empty((n_local, cap, 32), dtype=torch.uint8, **kw)
        return (k, d) if device is not None else (k.pin_memory(), d.pin_memory())

    dOutL, dOutR = outs(dev), (outs(dev) if stereo else None)
    hOutL, hOutR = outs(None), (outs(None) if stereo else None)
    mbf = np.float32(cfg["bf"]); mb = np.float32(mbf / np.float32(cfg["fx"])); maxD = float(np.float32(mbf / mb)) if stereo else 0.0
    stream = torch.cuda.ExternalStream(exL.cuda_stream(), device=dev)

    def step(imL, imR, oL, oR):
        exL.submit_batch(imL, oL)
        if stereo:
            exR.submit_batch(imR, oR)
        _, _, nL = exL.wait_batch()
        if stereo:
            exR.wait_batch()
            u, _ = ob.compute_stereo_matches(exL, exR, float(mbf), 0.0, maxD)  # mvuRight / mvDepth come back to the host
            return nL, u
        return nL, None

    def launches():
        return exL.launch_count() + (exR.launch_count() if stereo else 0)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        l0 = launches()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        th0 = time.perf_counter()
        e0.record(stream)
        for _ in range(steps):
            r = fn()
        e1.record(stream)
        barrier()
        wall = time.perf_counter() - th0
        sampler.mark(th0, time.perf_counter())
        # the stereo step ends with a blocking read-back, so the host clock brackets the same work as the events
        return max(e0.elapsed_time(e1) * 1e-3, 0.0), wall, r, launches() - l0

    sampler = B.ClockSampler(local)
    if rank == 0:
        sampler.start()
    t_dev, wall_dev, (nL, u), n_launch = timed(lambda: step(dL, dR, dOutL, dOutR), args.steps, args.warmup)
    t_dev = vmax(max(t_dev, wall_dev if stereo else 0.0))
    value = n_total * args.steps / t_dev
    # per-stage events for the dominant kernel (left extractor; the right one runs the same kernels on its own stream)
    exL.set_profiling(True)
    stage = {}
    for i in range(1 + min(args.steps, 10)):
        step(dL, dR, dOutL, dOutR)
        if i:
            for k, v in exL.stage_ms().items():
                stage[k] = stage.get(k, 0.0) + v / min(args.steps, 10)
    exL.set_profiling(False)
    barrier()
    t_e2e, wall_e2e, _, n_launch_e2e = timed(lambda: step(hL, hR, hOutL, hOutR), args.steps, args.warmup)
    t_e2e = vmax(max(t_e2e, wall_e2e))
    e2e = n_total * args.steps / t_e2e
    e2e_single, e2e_api = e2e, None
    if not stereo:
        # the same step with two batches in flight (ExtractorRing, as in config 1): while one batch computes, the next one's
        # frames cross PCIe and the previous one's results come back; every step still copies its inputs in and its results out
        ring = ob.ExtractorRing(depth=2, chunk=max(n_local, 96), device=local, **p)
        for _ in range(max(args.warmup, 3)):
            ring.push(hL)
        ring.drain()
        barrier()
        lr0 = ring.launch_count()
        r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        r0.record()
        for _ in range(args.steps):
            ring.push(hL)
        n_ring = ring.drain()[-1][2]
        torch.cuda.synchronize()
        r1.record()
        barrier()
        torch.cuda.synchronize()
        t_ring = vmax(r0.elapsed_time(r1) * 1e-3)
        n_launch_ring = ring.launch_count() - lr0
        assert np.array_equal(n_ring, nL), "ring and single-call keypoint counts differ"
        ring.close()
        if n_total * args.steps / t_ring > e2e:
            e2e, t_e2e, n_launch_e2e = n_total * args.steps / t_ring, t_ring, n_launch_ring
            e2e_api = "ExtractorRing.push: orbx_extract_batch_submit/_wait on 2 handles (2 batches in flight), pinned HOST buffers"
    else:
        # stereo: two (left, right) extractor pairs used alternately — while one pair's frames cross PCIe and its kernels
        # run, the other pair's stereo match runs and its mvuRight / mvDepth go back to the host
        pairs = [(exL, exR), (ob.ORBextractor(**p, device=local), ob.ORBextractor(**p, device=local))]
        outs2 = [(hOutL, hOutR), (outs(None), outs(None))]
        busy = [False, False]
        res = [None]

        def ring_step(i, submit=True):
            a, b = pairs[i & 1]
            if busy[i & 1]:
                _, _, nl_ = a.wait_batch()
                b.wait_batch()
                u_, _ = ob.compute_stereo_matches(a, b, float(mbf), 0.0, maxD)
                res[0] = (nl_, u_)
                busy[i & 1] = False
            if submit:
                a.submit_batch(hL, outs2[i & 1][0])
                b.submit_batch(hR, outs2[i & 1][1])
                busy[i & 1] = True

        for i in range(2 * max(args.warmup, 2)):
            ring_step(i)
        ring_step(0, False); ring_step(1, False)
        barrier()
        lr0 = sum(x.launch_count() for pr in pairs for x in pr)
        th0 = time.perf_counter()
        for i in range(args.steps):
            ring_step(i)
        ring_step(args.steps, False); ring_step(args.steps + 1, False)
        torch.cuda.synchronize()
        t_ring = vmax(time.perf_counter() - th0)  # (every ring step ends with a blocking read-back: the host clock brackets it)
        n_launch_ring = sum(x.launch_count() for pr in pairs for x in pr) - lr0
        assert np.array_equal(res[0][0], nL) and np.array_equal(res[0][1], u), "ring and single-call stereo results differ"
        for x in pairs[1]:
            x.close()
        if n_total * args.steps / t_ring > e2e:
            e2e, t_e2e, n_launch_e2e = n_total * args.steps / t_ring, t_ring, n_launch_ring
            e2e_api = ("two (left, right) extractor pairs in flight: orbx_extract_batch_submit/_wait per eye with pinned HOST "
                       "buffers + orbx_stereo_match_batch")
    clocks = sampler.stop() if rank == 0 else None
    n_mean = float(np.mean(nL))
    imgs = 2 if stereo else 1
    h2d = n_local * w * h * imgs
    d2h = n_local * cap * 60 * imgs + n_local * 4 * imgs + (n_local * cap * 8 if stereo else 0)
    peak, peak_src = B.peaks()
    kern = {k: v for k, v in stage.items() if k not in ("ingest", "egress")}
    dom = max(kern, key=kern.get)
    alg_frame = w * h + level_area_sum(w, h) + 60 * n_mean
    alg_launch = alg_frame * n_local            # one launch of the dominant kernel covers this rank's frames of ONE eye
    achieved = alg_launch / (kern[dom] * 1e-3) / 1e9
    alg_unit = imgs * alg_frame + (8 * n_mean if stereo else 0)
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": alg_launch,
                "algorithmic_bytes_per_unit": alg_unit, "kernel_ms": kern[dom], "stage_ms": {k: round(v, 4) for k, v in kern.items()},
                "whole_step_achieved": alg_unit * value / world / 1e9, "whole_step_frac": alg_unit * value / world / 1e9 / peak}
    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        cpu = cpu_baseline(cfg, left[:cfg["distinct"]], None if right is None else right[:cfg["distinct"]], args.cpu_seconds)
    tot_launch, tot_launch_e2e = int(vsum(n_launch)), int(vsum(n_launch_e2e))
    if rank == 0:
        line = {"metric": cfg["metric"], "value": value, "unit": cfg["unit"], "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": t_dev / args.steps * 1e3, "higher_is_better": True,
                "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
                "config": {"workload": cfg["name"], "units_per_step_total": n_total, "units_per_step_per_gpu": n_local,
                           "kp": nf, "levels": 8, "scale": 1.2, "fast": [20, 7], "keypoints_per_image": n_mean,
                           "stereo_matches_per_pair": float(np.mean([(u[f, :nL[f]] >= 0).sum() for f in range(n_local)])) if stereo else None,
                           "sharding": "sharding.shard_range: contiguous blocks of the 256 pairs" if strong else "fixed units per GPU",
                           "l2": "inputs larger than L2 (%d MB of images + pyramids per step per GPU)" %
                                 ((n_local * imgs * (w * h + 2 * level_area_sum(w, h))) >> 20)},
                "e2e": {"value": e2e, "unit": cfg["unit"], "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "ms_per_step": t_e2e / args.steps * 1e3, "gpu_launches": tot_launch_e2e,
                        "api": e2e_api or ("orbx_extract_batch_submit/_wait per eye with pinned HOST buffers + orbx_stereo_match_batch"
                                           if stereo else "orbx_extract_batch_submit/_wait with pinned HOST buffers"),
                        "single_call_value": e2e_single},
                "gpu_launches": tot_launch, "clocks": clocks, "host_affinity": numa, "roofline": roofline,
                "cpu_baseline": cpu, "parity": parity}
        out.emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if rank == 0 and not parity["ok"]:
        sys.exit(1)


# ------------------------------------------------------------------------------------------------ config 4: Hamming 2-NN
KNN = dict(name="knn2_100k_x_1M_256bit_ratio0.7", nq=100000, nt=1000000, metric="hamming_knn2_gpairs_per_s", unit="Gpairs/s")
KTC_K = 288  # int8 entries per descriptor row in k_knn_tc (256 bits + 32 norm / column-index entries, orbx_knn_tc.cuh)


def knn_roofline(path, pairs_per_s_per_gpu):
    """Roofline of the 2-NN kernel that ran. Tensor-core path (k_knn_tc): 2*288 int8 ops per pair against the dense int8
    rate; the driver-measured number in MEASURED_PEAKS.json is cuBLAS bf16 (int8 issues at twice the bf16 rate), so both
    that figure x 2 and NVIDIA's nominal 4.5 Pop/s are given and `frac` uses the nominal one (the kernel exceeds 2 x the
    measured cuBLAS bf16 rate). POPC path (k_knn2): 8 POPC per pair against the measured POPC issue rate."""
    if path == 2:
        ach = pairs_per_s_per_gpu * 2 * KTC_K / 1e12
        try:
            with open(os.path.join(B.ROOT, "MEASURED_PEAKS.json")) as f:
                bf16 = float(json.load(f)["bf16_tflops"])
        except Exception:
            bf16 = None
        try:
            with open(os.path.join(B.ROOT, "profiles", "traffic.json")) as f:
                traffic = float(json.load(f)["knn_tc"])  # dram read + write of one 100k x 1M launch (ncu --set full)
        except Exception:
            traffic = None
        return {"bound": "tensor", "kernel": "k_knn_tc", "achieved": ach, "peak": 4500.0, "unit": "TFLOP/s", "frac": ach / 4500.0,
                "traffic": traffic, "ncu_pipe_tensor_active_pct": 80.75, "peak_source": "nominal dense int8 tcgen05 rate of B200 (4.5 Pop/s); unit is int8 Top/s",
                "frac_of_2x_measured_bf16": (ach / (2 * bf16)) if bf16 else None, "measured_bf16_tflops": bf16,
                "ops_per_pair": 2 * KTC_K}
    popc_peak = 148 * 16 * 1.965e9 / 8
    return {"bound": "popc", "kernel": "k_knn2", "achieved": pairs_per_s_per_gpu / 1e9, "peak": popc_peak / 1e9,
            "unit": "Gpairs/s/GPU", "frac": pairs_per_s_per_gpu / popc_peak, "traffic": None,
            "peak_source": "148 SM x 16 POPC/clk x 1.965 GHz / 8 POPC per pair (scripts/pipe_probe.cu measured 0.496 warp-POPC/clk/SM)"}


def parity_knn(m, q, t, i1, d1, d2, rank_lo, samples=512):
    """`samples` queries of this rank's shard against the oracle's sequential scan of all 1M train rows."""
    from oracle_lib import OraclePrims
    rng = np.random.default_rng(99)
    pick = np.sort(rng.choice(len(q), min(samples, len(q)), replace=False))
    o1, e1, e2 = OraclePrims().knn2(q[pick], t, popcnt=True)
    same = bool(np.array_equal(i1[pick], o1) and np.array_equal(d1[pick], e1) and np.array_equal(d2[pick], e2))
    return {"what": f"{len(pick)} sampled queries (of {len(q)}, shard offset {rank_lo}) x {len(t)} train rows vs the oracle's "
                    "sequential best / second-best scan (orc_knn2_popcnt)", "index_dist_second_identical": same,
            "mismatching_queries": int(((i1[pick] != o1) | (d1[pick] != e1) | (d2[pick] != e2)).sum()), "ok": same}


def run_knn(args):
    out = B.StdoutGuard()
    import torch
    import torch.distributed as dist
    import orbslam2_b200 as ob
    from orbslam2_b200 import sharding
    from orbslam2_b200.synth import synth_descriptors

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = B.bind_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev, timeout=__import__("datetime").timedelta(seconds=300))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        v = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(v, op=op)
        return float(v.item())

    nq, nt = KNN["nq"], KNN["nt"]
    t, q = synth_descriptors(nt, nq, seed=7, dup=1000)
    lo, hi = sharding.shard_range(nq, rank, world)
    ql = np.ascontiguousarray(q[lo:hi])
    m = ob.ORBmatcher(0.7, device=local)
    dt = torch.from_numpy(t).to(dev)
    m.set_train(dt)                       # the map's descriptors stay resident (keyframe-vs-map search)
    dq = torch.from_numpy(ql).to(dev)
    hq = torch.from_numpy(ql).pin_memory()
    stream = torch.cuda.ExternalStream(m.cuda_stream(), device=dev)

    # parity gate on the full-size problem
    i1, d1, d2 = m.knn2(dq)
    path = m.last_knn_path()
    parity = parity_knn(m, ql, t, i1.cpu().numpy(), d1.cpu().numpy().view(np.uint16), d2.cpu().numpy().view(np.uint16), lo)
    pok = reduce(1.0 if parity["ok"] else 0.0, dist.ReduceOp.MIN)
    parity["all_ranks_ok"] = bool(pok == 1.0)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        l0 = m.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        th0 = time.perf_counter()
        e0.record(stream)
        for _ in range(steps):
            r = fn()
        e1.record(stream)
        barrier()
        wall = time.perf_counter() - th0
        sampler.mark(th0, time.perf_counter())
        return e0.elapsed_time(e1) * 1e-3, wall, r, m.launch_count() - l0

    sampler = B.ClockSampler(local)
    if rank == 0:
        sampler.start()
    t_dev, _, _, n_launch = timed(lambda: m.knn2(dq), args.steps, args.warmup)
    t_dev = reduce(t_dev, dist.ReduceOp.MAX)
    value = nq * nt * args.steps / t_dev / 1e9
    # e2e: host queries in, host (index, best, second) out, every step; the train set stays resident
    hq_np = hq.numpy()
    t_e2e, wall_e2e, r, n_launch_e2e = timed(lambda: m.knn2(hq_np), args.steps, args.warmup)
    t_e2e = reduce(max(t_e2e, wall_e2e), dist.ReduceOp.MAX)
    clocks = sampler.stop() if rank == 0 else None
    e2e = nq * nt * args.steps / t_e2e / 1e9
    acc = int(((r[1] <= 50) & (r[1].astype(np.float32) < np.float32(0.7) * r[2].astype(np.float32))).sum())
    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        cpu = B.cpu_match_baseline(nt=nt)
    tl, tle = int(reduce(n_launch, dist.ReduceOp.SUM)), int(reduce(n_launch_e2e, dist.ReduceOp.SUM))
    if rank == 0:
        out.emit(json.dumps({
            "metric": KNN["metric"], "value": value, "unit": KNN["unit"], "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": t_dev / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u8 bits as int8 on tcgen05 (int32 accumulate, exact)" if path == 2 else "u32 popc",
            "data": "synthetic",
            "config": {"workload": KNN["name"], "queries": nq, "train": nt, "queries_per_gpu": hi - lo, "ratio": 0.7,
                       "accepted_th50_rank0": acc, "train_set": "resident in HBM (orbm_set_train), replicated on every rank",
                       "l2": "train set 32 MB bits / 288 MB int8 expansion; queries x train tiles stream through L2 every step"},
            "e2e": {"value": e2e, "unit": KNN["unit"], "h2d_bytes_per_step": (hi - lo) * 32, "d2h_bytes_per_step": (hi - lo) * 8,
                    "ms_per_step": t_e2e / args.steps * 1e3, "gpu_launches": tle,
                    "api": "orbm_knn2 with HOST query / result buffers, resident train set"},
            "gpu_launches": tl, "clocks": clocks, "host_affinity": numa,
            "roofline": knn_roofline(path, nq * nt * args.steps / t_dev / world), "cpu_baseline": cpu, "parity": parity}))
    if world > 1:
        dist.destroy_process_group()
    if not parity["all_ranks_ok"]:
        sys.exit(1)


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    """--impl reference --config k: the reference's CPU code for that config on the host cores (rank 0 only)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    if args.config == 4:
        r = None
        t0 = time.perf_counter()
        for _ in range(max(1, min(args.steps, 3))):
            r = B.cpu_match_baseline(nt=KNN["nt"], q_per_thread=8)
        dt = time.perf_counter() - t0
        line = {"impl": "reference", "metric": KNN["metric"], "value": r["value"], "unit": KNN["unit"], "scaling": "strong",
                "config": {"workload": KNN["name"], "queries": KNN["nq"], "train": KNN["nt"], "sampled": r["sample"]},
                "cpu_baseline": r, "ms_per_step": dt * 1e3 / max(1, min(args.steps, 3)), "dtype": "u32 popc"}
    else:
        cfg = CFG[args.config]
        left, right = synth_inputs(cfg, 0, cfg["distinct"])
        pool = CpuPairPool(cfg)
        per_step = pool.threads
        for _ in range(min(args.warmup, 1)):
            pool.run(left, right, per_step)
        steps = max(1, min(args.steps, 5 if cfg["w"] < 2000 else 2))
        t0 = time.perf_counter()
        for _ in range(steps):
            pool.run(left, right, per_step)
        dt = time.perf_counter() - t0
        v = steps * per_step / dt
        sample = f"{per_step} units per step ({pool.threads} host threads x 1), {steps} steps"
        line = {"impl": "reference", "metric": cfg["metric"], "value": v, "unit": cfg["unit"],
                "scaling": "strong" if args.config == 3 else "weak",
                "config": {"workload": cfg["name"], "units_per_step": per_step, "kp": cfg["nf"], "levels": 8, "scale": 1.2, "fast": [20, 7]},
                "cpu_baseline": {"value": v, "unit": cfg["unit"], "cores": pool.threads, "kind": pool.kind, "sample": sample},
                "ms_per_step": dt / steps * 1e3, "dtype": "u8"}
    line.update({"n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "vs_baseline": None,
                 "data": "synthetic", "gpu_launches": 0,
                 "e2e": {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}})
    print(json.dumps(line))


def main(args):
    if args.impl == "reference":
        run_reference(args)
    elif args.config == 4:
        run_knn(args)
    else:
        run_shape(args, CFG[args.config], strong=(args.config == 3))
