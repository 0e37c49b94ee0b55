#!/usr/bin/env python
"""bench.py — headline benchmark of the ORB front-end hot path (BASELINE.json: "ORB extract frames/s @640x480/1000kp at
1/2/4/8 B200; Hamming match Gpairs/s").

    python bench.py --gpus N --steps K --warmup W            # our CUDA path (one process per GPU; torchrun for N>1)
    python bench.py --impl reference --gpus N --steps K ...   # the reference's own CPU ORBextractor on the host cores

One "step" = one pass of the extraction hot path (pyramid -> FAST -> quadtree -> orientation -> blur -> rBRIEF) over one
batch of synthetic 640x480 frames with the TUM1 settings (1000 kp, 8 levels, 1.2, FAST 20/7). Per-GPU work is fixed as
N grows (frames shard independently, no collective on the data path) => "scaling": "weak".

  value      frames/s with the batch already resident in HBM and results left in HBM (device pointers through the C ABI)
  e2e        frames/s through the C-ABI call with HOST buffers: pinned host images in, keypoints+descriptors out, both
             copies of every step inside the timed region. `value` keeps two batches in flight (ExtractorRing: the
             submit/wait halves of orbx_extract_batch on two handles); `single_call_value` is one blocking
             orbx_extract_batch per step
  roofline   dominant kernel: algorithmic bytes per launch (W*H + sum of level areas + 60*N_out per
             frame, times the frames in the launch) / its mean launch duration (CUDA events on the launching stream)
  cpu_baseline  the reference's ORBextractor.cc (oracle/_ref, compiled unmodified on the cv shim) on the host cores
  match      secondary metric: brute-force Hamming 2-NN Gpairs/s (config 4 shape, queries sharded across ranks)

Inputs are larger than L2: the default batch of 384 frames is 118 MB of images plus 2 x 370 MB of pyramid/blurred
levels per step, so nothing survives in the 126 MB L2 from one step to the next.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402

W, H, NFEAT = 640, 480, 1000
PARAMS = dict(nfeatures=NFEAT, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, patchSize=31, halfPatchSize=15,
              edgeThreshold=19)
METRIC = "orb_extract_frames_per_s_640x480_1000kp"
UNIT = "frames/s"
N_DISTINCT = 64  # distinct synthetic frames; a batch tiles them (content repeats, addresses do not)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="timed steps (default: 200 for config 1, 10 otherwise)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=1, choices=[1, 2, 3, 4, 5],
                    help="BASELINE.json config: 1 = 640x480/1000kp extraction (the headline, default), 2 = KITTI-shape stereo, "
                         "3 = EuRoC-shape stereo, 256 pairs sharded over the ranks (strong scaling), 4 = Hamming 2-NN 100k x 1M, "
                         "5 = 4K frames")
    ap.add_argument("--batch", type=int, default=384, help="frames per step per GPU")
    ap.add_argument("--chunk", type=int, default=0, help="frames per pipeline chunk of the host-buffer path (0 = default)")
    ap.add_argument("--dev-chunk", type=int, default=0, help="frames per chunk of the device-buffer path (0 = one launch sequence)")
    ap.add_argument("--ring-depth", type=int, default=2, help="batches in flight for the e2e number")
    ap.add_argument("--ring-chunk", type=int, default=192)  # measured 64 / 96 / 128 / 192: 165.8k / 167.6k / 167.6k / 169.4k frames/s
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-match", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline leg")
    a = ap.parse_args()
    if a.steps is None:
        a.steps = 200 if a.config == 1 and a.impl == "ours" else (20 if a.config == 1 else 10)
    return a


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel, batch):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel, from the committed
    `ncu --set full` capture of this same command (profiles/traffic.json: {"batch": B, "<stage>": bytes}); None when
    the capture was taken at another batch size."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        return float(t[kernel]) if int(t.get("batch", -1)) == batch and kernel in t else None
    except Exception:
        return None


def level_area_sum(w, h):
    """Sum of pyramid level areas, same rounding as the extractor (src/ORBextractor.cc:925-926)."""
    total, sf = 0, np.float32(1.0)
    for l in range(PARAMS["nlevels"]):
        inv = np.float32(1.0) / sf
        total += int(np.rint(np.float32(w) * inv)) * int(np.rint(np.float32(h) * inv))
        sf = np.float32(np.float64(sf) * np.float64(np.float32(PARAMS["scaleFactor"])))
    return total


def algorithmic_bytes_per_frame(n_out):
    return W * H + level_area_sum(W, H) + 60 * n_out


def distinct_frames():
    from orbslam2_b200.synth import synth_image
    return np.stack([synth_image(W, H, s) for s in range(N_DISTINCT)])


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []      # (arrival time, csv line)
        self.windows = []    # timed regions [t0, t1] (host clock); samples outside them are warm-up / idle

    def mark(self, t0, t1):
        self.windows.append((t0, t1))

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [ln for t, ln in self.lines if any(a <= t <= b + 0.05 for a, b in self.windows)]
        for ln in (inside if len(inside) >= 3 or not self.lines else [ln for _, ln in self.lines]):
            c = [x.strip() for x in ln.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "samples_in_timed_regions": len(inside)}


# ------------------------------------------------------------------------------------------------ CPU arm
try:
    ORIG_AFFINITY = os.sched_getaffinity(0)  # the GPU arm narrows the affinity to the GPU's NUMA node; CPU legs undo that
except Exception:
    ORIG_AFFINITY = None


def cpu_cores():
    try:
        if ORIG_AFFINITY is not None:
            os.sched_setaffinity(0, ORIG_AFFINITY)
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class CpuExtractorPool:
    """The reference's CPU ORBextractor on every host core: one extractor object per thread (ctypes releases the GIL).
    kind "reference" = oracle/_ref/libref_orbextractor.so (the UNMODIFIED /root/reference/src/ORBextractor.cc compiled
    on the cv shim); kind "port" = oracle/liborb_oracle.so when the prebuilt reference object is absent."""

    def __init__(self, threads=None):
        import oracle_lib
        self.threads = threads or cpu_cores()
        if os.path.exists(oracle_lib.REF_SO):
            self.kind, cls = "reference", oracle_lib.RefExtractor
        else:
            self.kind, cls = "port", oracle_lib.Oracle
        self.ex = [cls(**PARAMS) for _ in range(self.threads)]

    def run(self, frames, count):
        """Extract `count` frames (cycling through `frames`) split over the threads; returns (seconds, keypoints)."""
        nk = [0] * self.threads

        def work(t):
            for i in range(t, count, self.threads):
                k, _ = self.ex[t].extract(frames[i % len(frames)], 4 * NFEAT)
                nk[t] += len(k)

        th = [threading.Thread(target=work, args=(t,)) for t in range(self.threads)]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        return time.perf_counter() - t0, sum(nk)


CPU_ARM_NOTE = ("unmodified reference src/ORBextractor.cc, g++ -O3 -march=x86-64-v3 -ffp-contract=off, on the cv shim (OpenCV's "
                "resize / copyMakeBorder / FAST / GaussianBlur restated bit-exactly in oracle/cvprim.hpp with vectorisable loops); "
                "within 25 % of the cv2-hybrid baseline on real OpenCV 4.13 (profiles/r02_cv2_hybrid_cpu_baseline.json)")


def cpu_baseline(frames, seconds):
    pool = CpuExtractorPool()
    dt, _ = pool.run(frames, pool.threads)  # warm-up + rate estimate
    per_round = max(dt, 1e-3)
    rounds = int(max(1, min(64, (seconds - dt) / per_round)))
    count = pool.threads * rounds
    dt, nk = pool.run(frames, count)
    return {"value": count / dt, "unit": UNIT, "cores": pool.threads, "kind": pool.kind, "build": CPU_ARM_NOTE,
            "sample": f"{count} synthetic 640x480 frames, {pool.threads} threads x {rounds} frames each, "
                      f"{dt:.1f} s wall; {nk / count:.0f} keypoints/frame"}


def cpu_match_baseline(nt=1000000, q_per_thread=2500):
    """The reference's DescriptorDistance loop (SWAR popcount on 8 x int32, src/ORBmatcher.cc:1643-1659) scanning the
    whole train matrix with the best / second-best rule, one block of queries per host thread (the oracle's restatement of the
    scan, pinned to the unmodified ORBmatcher.cc by tests/test_cpu_refmatcher.py). A bounded sample of ~10-15 s:
    threads x q_per_thread queries x 1M train rows."""
    import oracle_lib
    from orbslam2_b200.synth import synth_descriptors
    threads = cpu_cores()
    t, q = synth_descriptors(nt, threads * q_per_thread, seed=7, dup=1000)
    P = [oracle_lib.OraclePrims() for _ in range(threads)]

    def work(k):
        P[k].knn2(q[k * q_per_thread:(k + 1) * q_per_thread], t)

    th = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
    t0 = time.perf_counter()
    for x in th:
        x.start()
    for x in th:
        x.join()
    dt = time.perf_counter() - t0
    pairs = threads * q_per_thread * nt
    return {"value": pairs / dt / 1e9, "unit": "Gpairs/s", "cores": threads, "kind": "port",
            "sample": f"{threads * q_per_thread} queries x {nt} train rows, {threads} threads, {dt:.1f} s wall; "
                      "reference SWAR DescriptorDistance + best/second-best scan"}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation on the host cores, same metric/config/unit."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    frames = distinct_frames()
    pool = CpuExtractorPool()
    per_step = pool.threads * 8  # bounded sample per step (8 frames per thread: ~0.1 s; with 2 the thread start / join tail cost 20-35 %)
    for _ in range(args.warmup):
        pool.run(frames, per_step)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        pool.run(frames, per_step)
    dt = time.perf_counter() - t0
    v = args.steps * per_step / dt
    sample = f"{per_step} frames per step ({pool.threads} host threads x 8), {args.steps} steps"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "orb_extract_640x480_1000kp_tum1", "frames_per_step": per_step, "kp": NFEAT,
                   "levels": 8, "scale": 1.2, "fast": [20, 7]},
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": pool.threads, "kind": pool.kind, "build": CPU_ARM_NOTE, "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def bind_to_gpu_numa_node(index):
    """Pin this rank's threads to the CPUs local to its GPU (NVML's ideal CPU affinity), so the pinned staging buffers
    it allocates next are first-touched on the GPU's NUMA node: with 8 ranks the host-to-device copies otherwise
    cross the socket interconnect and end-to-end throughput stops scaling. Returns a description for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        pynvml.nvmlDeviceSetCpuAffinity(h)
        return f"bound to {len(os.sched_getaffinity(0))} GPU-local CPUs"
    except Exception as e:  # affinity is an optimisation, never a requirement
        return f"not bound ({type(e).__name__})"


class StdoutGuard:
    """Libraries (NCCL's version banner) print to fd 1; the contract is ONE JSON line on stdout. Everything written to
    fd 1 while the guard is active goes to stderr; emit() writes the JSON line to the real stdout."""

    def __init__(self):
        sys.stdout.flush()
        self.real = os.dup(1)
        os.dup2(2, 1)

    def emit(self, line):
        sys.stdout.flush()
        os.write(self.real, (line + "\n").encode())


# ------------------------------------------------------------------------------------------------ GPU arm
def run_ours(args):
    out = StdoutGuard()
    import torch
    import torch.distributed as dist
    import orbslam2_b200 as ob
    from orbslam2_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(local)  # before any pinned allocation: first touch places the staging buffers
    if world > 1:
        dist.init_process_group("nccl", device_id=dev, timeout=__import__("datetime").timedelta(seconds=300))  # a stuck rank fails in minutes

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    B = args.batch
    frames = distinct_frames()
    # parity gate: one frame of the workload through the public call and through the CPU oracle, before timing
    import bench_configs
    cfg1 = dict(name="orb_extract_640x480_1000kp_tum1", w=W, h=H, nf=NFEAT, stereo=False)
    parity = bench_configs.parity_shape(cfg1, ob, local, frames[rank % N_DISTINCT][None], None) if rank == 0 else None
    # every rank gets its own shard of the synthetic frame sequence (different rotation of the distinct frames)
    idx = (np.arange(B) + rank * B) % N_DISTINCT
    host = torch.from_numpy(frames[idx]).pin_memory()          # [B,H,W] u8, pinned
    d_imgs = host.to(dev)
    ex = ob.ORBextractor(**PARAMS, device=local)
    cap = ex.max_keypoints(W, H)
    d_out = (torch.empty((B, cap, 7), dtype=torch.float32, device=dev),
             torch.empty((B, cap, 32), dtype=torch.uint8, device=dev))
    h_kps = torch.empty((B, cap, 7), dtype=torch.float32).pin_memory()
    h_desc = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
    h_n = np.zeros(B, np.int32)
    stream = torch.cuda.ExternalStream(ex.cuda_stream(), device=dev)

    def step_device():
        return ex.extract_batch(d_imgs, out=d_out)[2]

    def step_host():
        ex._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(ex._L.orbx_extract_batch(ex._h, host.data_ptr(), B, W, H, W, W * H, h_kps.data_ptr(),
                                            h_desc.data_ptr(), cap, h_n.ctypes.data))
        return h_n

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        l0 = ex.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        th0 = time.perf_counter()
        e0.record(stream)
        for _ in range(steps):
            n = fn()
        e1.record(stream)
        barrier()
        sampler.mark(th0, time.perf_counter())
        return e0.elapsed_time(e1) * 1e-3, n, ex.launch_count() - l0

    # ---- device-resident throughput (value)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ex.set_chunk(args.dev_chunk)
    t_dev, n, launches = timed(step_device, args.steps, args.warmup)
    # ---- per-stage CUDA events (on the launching stream) for the roofline of the dominant kernel: same steps, whole-batch launches
    ex.set_chunk(0)
    ex.set_profiling(True)
    step_device()
    stage_sum = {}
    for _ in range(args.steps):
        step_device()
        for k, v in ex.stage_ms().items():
            stage_sum[k] = stage_sum.get(k, 0.0) + v
    barrier()
    n_out_mean = float(np.mean(n))
    t_dev = max_over_ranks(t_dev)
    value = world * B * args.steps / t_dev
    ex.set_profiling(False)

    # ---- end to end through the C ABI with host buffers
    ex.set_chunk(args.chunk)
    t_e2e, _, launches_e2e = timed(step_host, args.steps, args.warmup)
    t_e2e = max_over_ranks(t_e2e)
    clocks = sampler.stop() if rank == 0 else None
    e2e = world * B * args.steps / t_e2e

    # ---- the same e2e step with two batches in flight (ExtractorRing: two handles, submit/wait halves of the C ABI call):
    # while one batch computes, the next one's frames cross PCIe and the previous one's results come back. Every step
    # still copies its own inputs in and its own results out inside the timed region.
    ring = ob.ExtractorRing(depth=args.ring_depth, chunk=args.ring_chunk, device=local, **PARAMS)
    for _ in range(max(args.warmup, args.ring_depth + 1)):
        ring.push(host)
    ring.drain()
    barrier()
    l0 = ring.launch_count()
    d0, d1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    d0.record()  # nothing is in flight: the event is stamped at once
    for _ in range(args.steps):
        ring.push(host)
    n_ring = ring.drain()[-1][2]
    torch.cuda.synchronize()
    d1.record()
    barrier()
    launches_ring = ring.launch_count() - l0
    assert np.array_equal(n_ring, h_n), "ring and single-call keypoint counts differ"
    t_ring = max_over_ranks(d0.elapsed_time(d1) * 1e-3)
    e2e_ring = world * B * args.steps / t_ring
    ring.close()

    # what the host link alone sustains with every rank copying at once (explains e2e at N > 1: the ranks share the
    # host's PCIe / memory fabric, so per-GPU copy bandwidth drops as N grows while the kernels scale linearly)
    cs = torch.cuda.Stream(device=dev)  # a torch-owned stream: the pinned-memory allocator keeps events on it
    with torch.cuda.stream(cs):
        for _ in range(2):
            d_imgs.copy_(host, non_blocking=True)
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record(cs)
        for _ in range(10):
            d_imgs.copy_(host, non_blocking=True)
        c1.record(cs)
        barrier()
    h2d_gbs = B * W * H * 10 / max_over_ranks(c0.elapsed_time(c1) * 1e-3) / 1e9
    h2d = B * W * H
    d2h = B * cap * (28 + 32) + B * 4 + 4

    # ---- roofline of the dominant kernel
    peak, peak_src = peaks()
    kern = {k: v / args.steps for k, v in stage_sum.items() if k not in ("ingest", "egress")}
    dom = max(kern, key=kern.get)
    alg = algorithmic_bytes_per_frame(n_out_mean) * B
    achieved = alg / (kern[dom] * 1e-3) / 1e9
    step_ms = sum(kern.values())
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": ncu_traffic(dom, B), "peak_source": peak_src,
                "limiter": "ALU-pipe + issue slots (packed 16-bit min/max, 2 warp-inst/clk/SM; profiles/r02final_k_fast_summary.txt: "
                           "pipe_alu 78.5 % active, issue 77.7 %, pipe_fma 20.7 %, DRAM < 5 % of peak) - the HBM fraction is reported as the "
                           "contract asks",
                "algorithmic_bytes_per_launch": alg, "kernel_ms": kern[dom],
                "stage_ms": {k: round(v, 4) for k, v in kern.items()},
                "whole_step_achieved": alg / (step_ms * 1e-3) / 1e9, "whole_step_frac": alg / (step_ms * 1e-3) / 1e9 / peak}

    match = None if args.no_match else bench_match(ob, torch, dev, rank, world, barrier, max_over_ranks)
    if match is not None and rank == 0 and not args.no_cpu_baseline:
        match["cpu_baseline"] = cpu_match_baseline()
    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        cpu = cpu_baseline(frames, args.cpu_seconds)
    total_launches = int(sum_over_ranks(launches))
    total_launches_ring = int(sum_over_ranks(launches_ring))  # collectives on every rank, never inside `if rank == 0`
    total_launches_e2e = int(sum_over_ranks(launches_e2e))
    if rank == 0:
        out.emit(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": t_dev / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": "orb_extract_640x480_1000kp_tum1", "frames_per_step_per_gpu": B, "kp": NFEAT,
                       "levels": 8, "scale": 1.2, "fast": [20, 7], "keypoints_per_frame": n_out_mean,
                       "l2": "inputs larger than L2 (batch images + pyramids ~ %d MB per step)" %
                             ((B * (W * H + 2 * level_area_sum(W, H))) >> 20)},
            # both ways of driving the C ABI with host buffers are measured; the headline is the faster one (on a box
            # whose host link is the limit, e.g. 8 ranks behind one PCIe root, the ring cannot add anything)
            "e2e": {"value": max(e2e_ring, e2e), "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": min(t_ring, t_e2e) / args.steps * 1e3,
                    "api": ("ExtractorRing.push: orbx_extract_batch_submit/_wait on %d handles (%d batches in flight), "
                            "host buffers, chunk %d" % (args.ring_depth, args.ring_depth, args.ring_chunk))
                           if e2e_ring >= e2e else "one blocking orbx_extract_batch per step, host buffers",
                    "ring_value": e2e_ring, "single_call_value": e2e,
                    "gpu_launches": total_launches_ring if e2e_ring >= e2e else total_launches_e2e,
                    "h2d_copy_only_gbs_per_gpu": h2d_gbs,
                    "h2d_copy_only_frames_per_s": world * h2d_gbs * 1e9 / (W * H)},
            "gpu_launches": total_launches, "clocks": clocks, "host_affinity": numa, "roofline": roofline, "cpu_baseline": cpu,
            "parity": parity, "match": match}))
    if world > 1:
        dist.destroy_process_group()


def bench_match(ob, torch, dev, rank, world, barrier, max_over_ranks, nq_total=100000, nt=1000000, reps=3):
    """Secondary metric (config 4): brute-force Hamming 2-NN, queries sharded across ranks, train set replicated."""
    from orbslam2_b200.synth import synth_descriptors
    t, q = synth_descriptors(nt, nq_total, seed=7, dup=1000)
    per = nq_total // world
    dq = torch.from_numpy(q[rank * per:(rank + 1) * per]).to(dev)
    dt = torch.from_numpy(t).to(dev)
    m = ob.ORBmatcher(0.7, device=dev.index)
    m.set_train(dt)
    stream = torch.cuda.ExternalStream(m.cuda_stream(), device=dev)
    m.knn2(dq)
    barrier()
    l0 = m.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        i1, d1, d2 = m.knn2(dq)
    e1.record(stream)
    barrier()
    sec = max_over_ranks(e0.elapsed_time(e1) * 1e-3) / reps
    pairs = per * world * nt
    acc = ((d1.int() <= 50) & (d1.float() < 0.7 * d2.float())).sum().item()
    import bench_configs
    ql = q[rank * per:(rank + 1) * per]
    parity = bench_configs.parity_knn(m, ql, t, i1.cpu().numpy(), d1.cpu().numpy().view(np.uint16),
                                      d2.cpu().numpy().view(np.uint16), rank * per) if rank == 0 else None
    return {"metric": "hamming_knn2_gpairs_per_s", "value": pairs / sec / 1e9, "unit": "Gpairs/s",
            "config": {"workload": "knn2_100k_x_1M_256bit", "queries": per * world, "train": nt, "ratio": 0.7},
            "ms": sec * 1e3, "gpu_launches": (m.launch_count() - l0) * world, "accepted_th50_rank0": int(acc),
            "roofline": bench_configs.knn_roofline(m.last_knn_path(), pairs / sec / world), "parity": parity}


if __name__ == "__main__":
    a = parse_args()
    if a.config != 1:
        import bench_configs
        bench_configs.main(a)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
