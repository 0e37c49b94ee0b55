"""CPU tests of the host-side logic: synthetic generators, bench accounting, and the multi-GPU sharding path run as two
gloo processes (world_size 2) with the CPU oracle standing in for the per-rank compute (tests may use the oracle)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from orbslam2_b200 import sharding  # noqa: E402
from orbslam2_b200.synth import synth_image, synth_stereo_pair, synth_descriptors  # noqa: E402


def test_synth_is_deterministic_and_textured():
    a, b = synth_image(320, 240, 5), synth_image(320, 240, 5)
    assert np.array_equal(a, b) and a.dtype == np.uint8 and a.shape == (240, 320)
    assert not np.array_equal(a, synth_image(320, 240, 6))
    assert a.std() > 20
    # low-contrast quadrant (bottom-right) exists so the minThFAST retry path fires
    assert a[120:, 160:].std() < 0.5 * a[:120, :160].std()
    l, r = synth_stereo_pair(320, 240, 1)
    assert l.shape == r.shape == (240, 320) and not np.array_equal(l, r)
    t, q = synth_descriptors(5000, 100, seed=1, dup=100)
    assert t.shape == (5000, 32) and q.shape == (100, 32) and t.dtype == np.uint8


def test_bench_accounting_matches_survey():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.level_area_sum(640, 480) == 950532                 # SURVEY.md §8 geometry table
    assert bench.algorithmic_bytes_per_frame(1000) == 1317732
    peak, src = bench.peaks()
    assert peak > 1000 and ("measured" in src or "fallback" in src)


def test_reference_arm_prints_one_contract_line():
    """bench.py --impl reference (the reference's CPU ORBextractor on the host cores) needs no GPU: one JSON line with the
    metric / unit of the product arm, impl, cpu_baseline and a zero-copy e2e block. The product arm on a box without a GPU
    must refuse loudly instead of falling back to the CPU."""
    import json
    import subprocess
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "orb_extract_frames_per_s_640x480_1000kp" and d["unit"] == "frames/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["frames_per_step"] == 8 * d["cpu_baseline"]["cores"]
    import torch
    if not torch.cuda.is_available():
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "0"],
                           capture_output=True, text=True, timeout=600, cwd=ROOT)
        assert r.returncode != 0 and "no CPU fallback" in (r.stderr + r.stdout)


def test_shard_range_partitions():
    for n in (0, 1, 7, 256, 1001):
        for world in (1, 2, 3, 8):
            cover = []
            for r in range(world):
                lo, hi = sharding.shard_range(n, r, world)
                assert 0 <= lo <= hi <= n
                cover += list(range(lo, hi))
            assert cover == list(range(n))
            sizes = sharding.shard_sizes(n, world)
            assert max(sizes) - min(sizes) <= 1 and sum(sizes) == n
    with pytest.raises(ValueError):
        sharding.shard_range(4, 2, 2)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from oracle_lib import Oracle, OraclePrims, KP_DTYPE, TUM1
    dist.init_process_group("gloo", rank=rank, world_size=world)
    p = dict(TUM1); p["nfeatures"] = 200
    orc = Oracle(**p)
    imgs = np.stack([synth_image(208, 160, s) for s in range(5)])
    cap = 400

    def extract_fn(block):
        k = np.zeros((len(block), cap), KP_DTYPE); d = np.zeros((len(block), cap, 32), np.uint8)
        n = np.zeros(len(block), np.int32)
        for i, im in enumerate(block):
            kk, dd = orc.extract(im, cap)
            n[i] = len(kk); k[i, :n[i]] = kk; d[i, :n[i]] = dd
        return k, d, n

    lo, hi, local = sharding.extract_sharded(extract_fn, imgs, rank, world)
    k, d, n = sharding.gather_frame_results(local, len(imgs), cap, KP_DTYPE)
    # query sharding of the 2-NN search, train replicated
    t, q = synth_descriptors(2000, 101, seed=3, dup=50)
    P = OraclePrims()
    qlo, qhi, res = sharding.knn2_sharded(lambda qq: P.knn2(qq, t), q, rank, world)
    parts = [None] * world
    dist.all_gather_object(parts, (qlo, qhi, res))
    i1 = np.concatenate([r[2][0] for r in parts]); d1 = np.concatenate([r[2][1] for r in parts])
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), k=k, d=d, n=n, lo=lo, hi=hi, i1=i1, d1=d1)
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo_sharding(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    assert (int(r0["lo"]), int(r0["hi"]), int(r1["lo"]), int(r1["hi"])) == (0, 3, 3, 5)
    for key in ("k", "d", "n", "i1", "d1"):
        assert r0[key].tobytes() == r1[key].tobytes(), key  # every rank ends with the same gathered result
    # and it equals the unsharded run
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_lib import Oracle, OraclePrims, TUM1
    p = dict(TUM1); p["nfeatures"] = 200
    orc = Oracle(**p)
    for f in range(5):
        kk, dd = orc.extract(synth_image(208, 160, f), 400)
        assert r0["n"][f] == len(kk)
        assert r0["k"][f, :len(kk)].tobytes() == kk.tobytes() and np.array_equal(r0["d"][f, :len(kk)], dd)
    t, q = synth_descriptors(2000, 101, seed=3, dup=50)
    i1, d1, _ = OraclePrims().knn2(q, t)
    assert np.array_equal(r0["i1"], i1) and np.array_equal(r0["d1"], d1)
