#!/usr/bin/env python
"""Secondary benchmark lines for the other BASELINE.json configs (bench.py carries the headline, config 1 + config 4):

  config 2  KITTI-shape stereo 1241x376, 2000 kp: extraction of both images + ComputeStereoMatches, pairs/s
  config 3  EuRoC-shape stereo 752x480, 1200 kp, 256 pairs sharded over the ranks, pairs/s
  config 5  4K 3840x2160, 8000 kp frame batches, frames/s

    python scripts/bench_configs.py                       # one GPU
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/bench_configs.py

Inputs are resident in HBM when the timed region starts (torch CUDA tensors through the C ABI's device-pointer mode);
mvuRight / mvDepth come back to the host (they are the product of the stereo step). Timing: CUDA events on the left
extractor's stream, max over ranks. Each line also carries the fraction of the HBM roofline (SURVEY.md §8(d) bytes) and
the reference's CPU ORBextractor (oracle/_ref) on the host cores for the same shape (extraction only; rank 0).
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import numpy as np  # noqa: E402


def level_area_sum(w, h, nlevels=8, scale=1.2):
    total, sf = 0, np.float32(1.0)
    for _ in range(nlevels):
        inv = np.float32(1.0) / sf
        total += int(np.rint(np.float32(w) * inv)) * int(np.rint(np.float32(h) * inv))
        sf = np.float32(np.float64(sf) * np.float64(np.float32(scale)))
    return total


def main():
    import torch
    import torch.distributed as dist
    import orbslam2_b200 as ob
    from orbslam2_b200.synth import synth_image, synth_stereo_pair
    from oracle_lib import TUM1
    import bench as B

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    peak, peak_src = B.peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def vmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def cpu_ref(w, h, nf, seconds=6.0):
        if rank != 0:
            return None
        params = dict(B.PARAMS); params["nfeatures"] = nf
        saved = (B.PARAMS, B.NFEAT)
        B.PARAMS, B.NFEAT = params, nf
        try:
            frames = np.stack([synth_image(w, h, s) for s in range(4)])
            r = B.cpu_baseline(frames, seconds)
            r["sample"] = r["sample"].replace("640x480", f"{w}x{h}")
            return r
        finally:
            B.PARAMS, B.NFEAT = saved

    def stereo_config(name, w, h, nf, bf, fx, pairs_per_gpu, steps=10, warmup=3, distinct=4):
        p = dict(TUM1); p["nfeatures"] = nf
        exL, exR = ob.ORBextractor(**p, device=local), ob.ORBextractor(**p, device=local)
        ps = [synth_stereo_pair(w, h, s + rank * distinct) for s in range(distinct)]
        idx = np.arange(pairs_per_gpu) % distinct
        dL = torch.from_numpy(np.ascontiguousarray(np.stack([ps[i][0] for i in idx]))).to(dev)
        dR = torch.from_numpy(np.ascontiguousarray(np.stack([ps[i][1] for i in idx]))).to(dev)
        cap = exL.max_keypoints(w, h)
        outL = (torch.empty((pairs_per_gpu, cap, 7), dtype=torch.float32, device=dev),
                torch.empty((pairs_per_gpu, cap, 32), dtype=torch.uint8, device=dev))
        outR = (torch.empty_like(outL[0]), torch.empty_like(outL[1]))
        stream = torch.cuda.ExternalStream(exL.cuda_stream(), device=dev)

        def step():
            exL.submit_batch(dL, outL)   # both extractions in flight: the two handles' kernels interleave
            exR.submit_batch(dR, outR)
            _, _, nL = exL.wait_batch()
            exR.wait_batch()
            u, z = ob.compute_stereo_matches(exL, exR, bf, 0.0, fx)
            return nL, u

        for _ in range(warmup):
            step()
        barrier()
        l0 = exL.launch_count() + exR.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        t0 = time.perf_counter()
        for _ in range(steps):
            nL, u = step()
        e1.record(stream)
        barrier()
        wall = time.perf_counter() - t0
        sec = vmax(max(e0.elapsed_time(e1) * 1e-3, 0.0))
        n_mean = float(np.mean(nL))
        matched = float(np.mean([(u[f, :nL[f]] >= 0).sum() for f in range(len(nL))]))
        alg = 2 * (w * h + level_area_sum(w, h) + 60 * n_mean) + 8 * n_mean
        value = world * pairs_per_gpu * steps / sec
        if rank == 0:
            print(json.dumps({"metric": f"orb_stereo_pairs_per_s_{w}x{h}_{nf}kp", "config": name, "value": value,
                              "unit": "pairs/s", "n_gpus": world, "pairs_per_step_per_gpu": pairs_per_gpu, "steps": steps,
                              "ms_per_step": sec / steps * 1e3, "host_wall_ms_per_step": wall / steps * 1e3,
                              "keypoints_per_image": n_mean, "stereo_matches_per_pair": matched,
                              "gpu_launches": exL.launch_count() + exR.launch_count() - l0,
                              "roofline": {"bound": "hbm", "achieved": alg * value / world / 1e9, "peak": peak,
                                           "unit": "GB/s", "frac": alg * value / world / 1e9 / peak,
                                           "algorithmic_bytes_per_pair": alg, "peak_source": peak_src},
                              "cpu_baseline_extract_only": cpu_ref(w, h, nf)}), flush=True)
        exL.close(); exR.close()

    def mono_config(name, w, h, nf, frames_per_gpu, steps=10, warmup=3, distinct=2):
        p = dict(TUM1); p["nfeatures"] = nf
        ex = ob.ORBextractor(**p, device=local)
        fr = [synth_image(w, h, s + rank * distinct) for s in range(distinct)]
        d = torch.from_numpy(np.stack([fr[i % distinct] for i in range(frames_per_gpu)])).to(dev)
        cap = ex.max_keypoints(w, h)
        out = (torch.empty((frames_per_gpu, cap, 7), dtype=torch.float32, device=dev),
               torch.empty((frames_per_gpu, cap, 32), dtype=torch.uint8, device=dev))
        stream = torch.cuda.ExternalStream(ex.cuda_stream(), device=dev)
        for _ in range(warmup):
            ex.extract_batch(d, out=out)
        barrier()
        l0 = ex.launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            _, _, n = ex.extract_batch(d, out=out)
        e1.record(stream)
        barrier()
        sec = vmax(e0.elapsed_time(e1) * 1e-3)
        n_mean = float(np.mean(n))
        alg = w * h + level_area_sum(w, h) + 60 * n_mean
        value = world * frames_per_gpu * steps / sec
        if rank == 0:
            print(json.dumps({"metric": f"orb_extract_frames_per_s_{w}x{h}_{nf}kp", "config": name, "value": value,
                              "unit": "frames/s", "n_gpus": world, "frames_per_step_per_gpu": frames_per_gpu,
                              "steps": steps, "ms_per_step": sec / steps * 1e3, "keypoints_per_frame": n_mean,
                              "gpu_launches": ex.launch_count() - l0,
                              "roofline": {"bound": "hbm", "achieved": alg * value / world / 1e9, "peak": peak,
                                           "unit": "GB/s", "frac": alg * value / world / 1e9 / peak,
                                           "algorithmic_bytes_per_frame": alg, "peak_source": peak_src},
                              "cpu_baseline": cpu_ref(w, h, nf)}), flush=True)
        ex.close()

    stereo_config("config2_kitti_stereo", 1241, 376, 2000, 386.1448, 718.856, 64)
    stereo_config("config3_euroc_stereo_256_pairs_sharded", 752, 480, 1200, 47.90639384423901, 435.2046959714599,
                  256 // world)
    mono_config("config5_4k", 3840, 2160, 8000, 16)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
