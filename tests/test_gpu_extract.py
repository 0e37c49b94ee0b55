"""GPU parity: liborbx (through the C ABI / Python mirror) vs the CPU oracle, stage by stage.

Bars (BASELINE.json north_star): pyramid + blurred pixels, FAST candidates/scores and the selected keypoint set
bit-exact; coordinates within 1e-4 px, angles within 1e-2 deg, >= 99.99 % of descriptors bit-identical."""
import numpy as np
import pytest

from oracle_lib import Oracle, canonical, TUM1
from orbslam2_b200.synth import synth_image

pytestmark = pytest.mark.gpu

SHAPES = [(640, 480, 1000), (752, 480, 1200), (1241, 376, 2000)]


@pytest.fixture(scope="module")
def ob():
    import orbslam2_b200
    return orbslam2_b200


def _params(nf):
    p = dict(TUM1); p["nfeatures"] = nf
    return p


def _compare_frame(ex, orc, img, frame=0, desc_stats=None, kps=None, desc=None):
    p = orc.params
    ko, do = orc.extract(img, 4 * p["nfeatures"])
    for l in range(p["nlevels"]):
        assert ex.level_size(l) == orc.level_size(l)
        assert np.array_equal(ex.pyramid_level(l, frame, True), orc.level(l, True)), f"pyramid level {l}"
        assert np.array_equal(ex.blurred_level(l, frame), orc.blurred(l)), f"blurred level {l}"
        cg = ex.candidates(l, frame); co = orc.candidates(l)
        cg = cg[np.lexsort((cg[:, 0], cg[:, 1]))]; co = co[np.lexsort((co[:, 0], co[:, 1]))]
        assert np.array_equal(cg, co), f"FAST candidates level {l}: {len(cg)} vs {len(co)}"
    ko, do = canonical(ko, do)
    assert len(kps) == len(ko)
    # GPU output is already canonical (level, y, x)
    assert np.array_equal(kps["octave"], ko["octave"])
    assert np.array_equal(kps["response"], ko["response"]) and np.array_equal(kps["size"], ko["size"])
    assert np.array_equal(kps["class_id"], ko["class_id"])
    assert np.abs(kps["x"] - ko["x"]).max() <= 1e-4 and np.abs(kps["y"] - ko["y"]).max() <= 1e-4
    assert np.abs(kps["angle"] - ko["angle"]).max() <= 1e-2
    same = (desc == do).all(axis=1)
    if desc_stats is not None:
        desc_stats[0] += int(same.sum()); desc_stats[1] += len(same)
    return same


@pytest.mark.parametrize("w,h,nf", SHAPES)
def test_single_frame_all_stages(ob, w, h, nf):
    p = _params(nf)
    ex = ob.ORBextractor(**p); orc = Oracle(**p)
    stats = [0, 0]
    for seed in (0, 1):
        img = synth_image(w, h, seed)
        kps, desc = ex(img)
        _compare_frame(ex, orc, img, 0, stats, kps, desc)
    assert stats[0] >= 0.9999 * stats[1], f"descriptors identical {stats[0]}/{stats[1]}"
    ex.close()


def test_batch_matches_single(ob):
    p = _params(1000)
    ex = ob.ORBextractor(**p); orc = Oracle(**p)
    imgs = np.stack([synth_image(640, 480, s) for s in range(6)])
    kps, desc, n = ex.extract_batch(imgs)
    stats = [0, 0]
    for f in (0, 3, 5):
        _compare_frame(ex, orc, imgs[f], f, stats, kps[f, :n[f]], desc[f, :n[f]])
    assert stats[0] >= 0.9999 * stats[1]
    # the same frames one at a time give identical bytes
    for f in range(6):
        k1, d1 = ex(imgs[f])
        assert np.array_equal(k1, kps[f, :n[f]]) and np.array_equal(d1, desc[f, :n[f]])
    ex.close()


def test_strided_input_and_noise(ob):
    """Row-strided host input (an ROI view) and a dense-corner noise image (candidate-capacity stress)."""
    p = _params(500)
    ex = ob.ORBextractor(**p); orc = Oracle(**p)
    big = np.random.default_rng(5).integers(0, 256, (300, 500), dtype=np.uint8)
    view = big[10:270, 20:420]
    k, d = ex(view[None][0])
    kb, db, nb = ex.extract_batch(view[None])
    img = np.ascontiguousarray(view)
    _compare_frame(ex, orc, img, 0, None, kb[0, :nb[0]], db[0, :nb[0]])
    assert np.array_equal(k, kb[0, :nb[0]])
    ex.close()


def test_flat_and_empty_images(ob):
    p = _params(1000)
    ex = ob.ORBextractor(**p)
    k, d = ex(np.full((480, 640), 77, np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)
    k, d = ex(np.zeros((0, 0), np.uint8))
    assert len(k) == 0
    with pytest.raises(ob.OrbxError):
        ex(np.zeros((60, 60), np.uint8))  # top pyramid levels smaller than one FAST cell
    ex.close()


def test_getters_match_oracle(ob):
    p = _params(1000)
    ex = ob.ORBextractor(**p); orc = Oracle(**p)
    t, pl, um = orc.tables()
    assert np.array_equal(ex.GetScaleFactors(), t[0]) and np.array_equal(ex.GetInverseScaleFactors(), t[1])
    assert np.array_equal(ex.GetScaleSigmaSquares(), t[2]) and np.array_equal(ex.GetInverseScaleSigmaSquares(), t[3])
    assert np.array_equal(ex.mnFeaturesPerLevel, pl) and np.array_equal(ex.umax, um)
    assert ex.GetLevels() == 8
    ex.close()


def test_device_io_torch(ob):
    import torch
    p = _params(1000)
    ex = ob.ORBextractor(**p)
    imgs = np.stack([synth_image(640, 480, s) for s in range(4)])
    kh, dh, nh = ex.extract_batch(imgs)
    kd, dd, nd = ex.extract_batch(torch.from_numpy(imgs).cuda())
    assert np.array_equal(nh, nd)
    for f in range(4):
        assert np.array_equal(dd[f, :nd[f]].cpu().numpy(), dh[f, :nh[f]])
        assert kd[f, :nd[f]].cpu().numpy().tobytes() == kh[f, :nh[f]].tobytes()
    ex.close()


def test_host_pipeline_chunks_match_single_launch(ob):
    """Host buffers are processed in software-pipelined chunks (copy-in / compute / copy-out overlap): every chunking,
    including a ragged last chunk, gives the bytes of the one-launch path."""
    p = _params(1000)
    ex = ob.ORBextractor(**p)
    imgs = np.stack([synth_image(640, 480, s % 5) for s in range(23)])
    ex.set_chunk(64)                      # 23 < 2*64 -> one launch sequence
    k0, d0, n0 = ex.extract_batch(imgs)
    for chunk in (1, 4, 7, 11):
        ex.set_chunk(chunk)
        k, d, n = ex.extract_batch(imgs)
        assert np.array_equal(n, n0)
        for f in range(len(imgs)):
            assert k[f, :n[f]].tobytes() == k0[f, :n0[f]].tobytes() and np.array_equal(d[f, :n[f]], d0[f, :n0[f]])
    # resident state (pyramid of a frame from a late chunk) is still addressable by absolute frame index
    lvl = ex.pyramid_level(3, 22)
    k1, d1 = ex(imgs[22])
    assert k1.tobytes() == k0[22, :n0[22]].tobytes() and np.array_equal(ex.pyramid_level(3, 0), lvl)
    ex.close()


def test_extractor_ring_matches_blocking_call(ob):
    """submit/wait halves of orbx_extract_batch on two handles (ExtractorRing): batches of different content pushed back
    to back come out in order with the bytes of the blocking call; misuse of the halves is an error, not a hang."""
    p = _params(1000)
    ex = ob.ORBextractor(**p)
    batches = [np.stack([synth_image(640, 480, (3 * b + s) % 7) for s in range(9)]) for b in range(5)]
    ref = [tuple(a.copy() for a in ex.extract_batch(b)) for b in batches]
    ring = ob.ExtractorRing(depth=2, chunk=4, **p)
    got = []
    for b in batches:
        r = ring.push(b)
        if r is not None:
            got.append(tuple(a.copy() for a in r))
    got += [tuple(a.copy() for a in r) for r in ring.drain()]
    assert len(got) == len(ref) and ring.drain() == []
    for (k, d, n), (k0, d0, n0) in zip(got, ref):
        assert np.array_equal(n, n0)
        for f in range(len(n)):
            assert k[f, :n[f]].tobytes() == k0[f, :n0[f]].tobytes() and np.array_equal(d[f, :n[f]], d0[f, :n0[f]])
    ring.close()
    # one pending batch per handle; wait without submit
    cap = ex.max_keypoints(640, 480)
    out = (np.zeros((9, cap), ob.KP_DTYPE), np.zeros((9, cap, 32), np.uint8))
    ex._pending = (None, out, np.zeros(9, np.int32))
    with pytest.raises(ob.OrbxError):
        ex.wait_batch()
    ex.submit_batch(batches[0], out)
    with pytest.raises(ob.OrbxError):
        ex.submit_batch(batches[1], out)
    k, d, n = ex.wait_batch()
    assert np.array_equal(n, ref[0][2]) and np.array_equal(d[0, :n[0]], ref[0][1][0, :n[0]])
    ex.close()


def test_device_submit_wait_on_two_handles(ob):
    """The halves with DEVICE buffers (torch CUDA tensors): left and right extractors of a stereo frame in flight
    together (the reference runs them on two host threads, src/Frame.cc:75-78) return the bytes of the blocking call."""
    import torch
    p = _params(1000)
    exL, exR = ob.ORBextractor(**p), ob.ORBextractor(**p)
    L = np.stack([synth_image(640, 480, s) for s in range(6)])
    R = np.stack([synth_image(640, 480, 10 + s) for s in range(6)])
    refL, refR = exL.extract_batch(L), exR.extract_batch(R)
    cap = exL.max_keypoints(640, 480)
    dL, dR = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    outs = [(torch.zeros((6, cap, 7), dtype=torch.float32, device="cuda"),
             torch.zeros((6, cap, 32), dtype=torch.uint8, device="cuda")) for _ in range(2)]
    for _ in range(2):  # twice: the second round reuses every buffer
        exL.submit_batch(dL, outs[0])
        exR.submit_batch(dR, outs[1])
        gotL, gotR = exL.wait_batch(), exR.wait_batch()
        for (k, d, n), (k0, d0, n0) in ((gotL, refL), (gotR, refR)):
            assert np.array_equal(n, n0)
            k, d = k.cpu().numpy(), d.cpu().numpy()
            for f in range(6):
                assert k[f, :n[f]].tobytes() == k0[f, :n0[f]].tobytes() and np.array_equal(d[f, :n[f]], d0[f, :n0[f]])
    exL.close(); exR.close()


def test_4k_frame_all_stages(ob):
    """BASELINE config 5: 3840x2160, 8000 keypoints — every stage bit-exact at the largest shape (26 925 FAST cells,
    per-level quotas up to 1737, TMA tiles at the widest cell size)."""
    p = _params(8000)
    ex = ob.ORBextractor(**p); orc = Oracle(**p)
    img = synth_image(3840, 2160, 0)
    kps, desc = ex(img)
    same = _compare_frame(ex, orc, img, 0, None, kps, desc)
    assert same.all(), f"descriptors identical {int(same.sum())}/{len(same)}"
    # a batch of two 4K frames through the chunked host path gives the same bytes
    ex.set_chunk(1)
    kb, db, nb = ex.extract_batch(np.stack([img, img[::-1].copy()]))
    assert kb[0, :nb[0]].tobytes() == kps.tobytes() and np.array_equal(db[0, :nb[0]], desc)
    ex.close()


# ---- wider inputs: many seeds per shape (final outputs), portrait, odd widths that move the 16-byte
# TMA realignment of k_fast, a pyramid whose top level is a single row of FAST cells, and low-contrast frames in which
# (almost) every cell takes the minThFAST retry
def _final_outputs_equal(ex, orc, img, stats):
    p = orc.params
    ko, do = canonical(*orc.extract(img, 4 * p["nfeatures"]))
    kps, desc = ex(img)
    assert len(kps) == len(ko), f"{len(kps)} vs {len(ko)} keypoints"
    for f in ("octave", "response", "size", "class_id"):
        assert np.array_equal(kps[f], ko[f]), f
    assert np.abs(kps["x"] - ko["x"]).max() <= 1e-4 and np.abs(kps["y"] - ko["y"]).max() <= 1e-4
    assert np.abs(kps["angle"] - ko["angle"]).max() <= 1e-2
    same = (desc == do).all(axis=1)
    stats[0] += int(same.sum()); stats[1] += len(same)


@pytest.mark.parametrize("w,h,nf,seeds", [(640, 480, 1000, 20), (752, 480, 1200, 20), (1241, 376, 2000, 20), (480, 640, 1000, 8),
                                          (641, 479, 1000, 6), (757, 483, 1200, 6), (643, 481, 1000, 4), (650, 490, 700, 4)])
def test_many_seeds_and_odd_shapes(ob, w, h, nf, seeds):
    p = _params(nf)
    ex = ob.ORBextractor(**p); orc = Oracle(**p)
    stats = [0, 0]
    for seed in range(100, 100 + seeds):
        _final_outputs_equal(ex, orc, synth_image(w, h, seed), stats)
    assert stats[0] >= 0.9999 * stats[1], f"descriptors identical {stats[0]}/{stats[1]}"
    ex.close()


def test_single_cell_row_top_level(ob):
    """Height chosen so that the top pyramid level holds exactly one row of FAST cells (nRows = 1)."""
    p = _params(600)
    orc = Oracle(**p)
    h = None
    for cand in range(150, 260):
        orc.extract(synth_image(640, cand, 3), 2400)
        hh = orc.level_size(7)[1]
        if (hh - 2 * 16) // 30 == 1:
            h = cand
            break
    assert h is not None
    ex = ob.ORBextractor(**p)
    stats = [0, 0]
    for seed in (3, 4):
        img = synth_image(640, h, seed)
        kps, desc = ex(img)
        _compare_frame(ex, orc, img, 0, stats, kps, desc)
    assert stats[0] >= 0.9999 * stats[1]
    ex.close()


@pytest.mark.parametrize("contrast", [0.15, 0.3])
def test_low_contrast_retry_cells(ob, contrast):
    """Contrast scaled down so that FAST at iniThFAST = 20 finds (almost) nothing and the cells fall back to minThFAST."""
    p = _params(1000)
    ex = ob.ORBextractor(**p); orc = Oracle(**p)
    stats = [0, 0]
    for seed in (0, 1, 2):
        img = synth_image(640, 480, seed).astype(np.float32)
        img = np.clip((img - 128.0) * contrast + 128.0, 0, 255).astype(np.uint8)
        kps, desc = ex(img)
        _compare_frame(ex, orc, img, 0, stats, kps, desc)
        ini = sum(int((orc.candidates(l)[:, 2] >= 20).sum()) for l in range(8))
        tot = sum(len(orc.candidates(l)) for l in range(8))
        assert tot > 500 and (contrast > 0.2 or ini < 0.2 * tot), (ini, tot)
    assert stats[0] >= 0.9999 * stats[1]
    ex.close()


def test_two_pass_fast_kernel_gives_identical_candidates(ob, monkeypatch):
    """k_fast2 (ORBX_FAST_V2=1: iniThFAST pre-test + work lists, the reference's two-threshold structure) against the
    default k_fast and the oracle: candidates, scores and final outputs identical (kept as a measured alternative,
    profiles/r02d_fast2_analysis.md)."""
    monkeypatch.setenv("ORBX_FAST_V2", "1")
    p = _params(1000)
    ex2 = ob.ORBextractor(**p)
    monkeypatch.delenv("ORBX_FAST_V2")
    ex1 = ob.ORBextractor(**p)
    orc = Oracle(**p)
    for seed, contrast in ((7, 1.0), (8, 0.2)):
        img = synth_image(640, 480, seed).astype(np.float32)
        img = np.clip((img - 128.0) * contrast + 128.0, 0, 255).astype(np.uint8)
        k2, d2 = ex2(img); k1, d1 = ex1(img)
        assert k1.tobytes() == k2.tobytes() and np.array_equal(d1, d2)
        _compare_frame(ex2, orc, img, 0, None, k2, d2)
    ex1.close(); ex2.close()


@pytest.mark.parametrize("nf", ["0", "1"])
def test_fp16_lane_fast_kernel_gives_identical_candidates(ob, monkeypatch, nf):
    """k_fast_t<3> (ORBX_FAST=3: tile expanded to fp16 lanes 1024 + p, circle operands by aligned LDS, part of the min/max
    tree as HFMA2.RELU pairs on the FMA pipe; ORBX_FAST3_NF=1 keeps the whole tree on the ALU pipe) against the default
    kernel and the oracle on shapes with different supertile geometry (profiles/r02x_fast3_analysis.md)."""
    monkeypatch.setenv("ORBX_FAST", "3")
    monkeypatch.setenv("ORBX_FAST3_NF", nf)
    p = _params(1000)
    ex3 = ob.ORBextractor(**p)
    monkeypatch.delenv("ORBX_FAST")
    monkeypatch.delenv("ORBX_FAST3_NF")
    ex1 = ob.ORBextractor(**p)
    orc = Oracle(**p)
    for (w, h), seed, contrast in (((640, 480), 7, 1.0), ((640, 480), 8, 0.2), ((757, 483), 9, 1.0), ((480, 640), 10, 1.0),
                                   ((1241, 376), 11, 0.5)):
        img = synth_image(w, h, seed).astype(np.float32)
        img = np.clip((img - 128.0) * contrast + 128.0, 0, 255).astype(np.uint8)
        k3, d3 = ex3(img); k1, d1 = ex1(img)
        assert k1.tobytes() == k3.tobytes() and np.array_equal(d1, d3)
        _compare_frame(ex3, orc, img, 0, None, k3, d3)
    ex1.close(); ex3.close()


@pytest.mark.parametrize("sf", [1.5, 1.9, 2.0, 2.5, 3.0])
def test_other_scale_factors(ob, sf):
    """The pyramid tables are packed for k_resize_fast (PRMT selectors in the x entries, both clamped source rows in the y
    entries); the generic k_resize reads the same tables. Scale factors other than 1.2 take every branch: 1.5 / 1.9 the fast
    kernel with other spans, 2.0 OpenCV's exact-2x shortcut, 2.5 / 3.0 the generic kernel (4 output pixels span > 8 source
    bytes). Everything against the oracle, as for the default configuration."""
    p = _params(500); p["scaleFactor"] = sf; p["nlevels"] = 3 if sf < 3 else 2  # (a third level at 3.0 is smaller than one FAST cell)
    ex = ob.ORBextractor(**p); orc = Oracle(**p)
    stats = [0, 0]
    for (w, h), seed in (((640, 480), 3), ((757, 483), 4)):
        img = synth_image(w, h, seed)
        kps, desc = ex(img)
        _compare_frame(ex, orc, img, 0, stats, kps, desc)
    imgs = np.stack([synth_image(640, 480, 20 + s) for s in range(5)])  # batch plan (72-row blur strips)
    kb, db, nb = ex.extract_batch(imgs)
    _compare_frame(ex, orc, imgs[4], 4, stats, kb[4, :nb[4]], db[4, :nb[4]])
    assert stats[0] >= 0.9999 * stats[1]
    ex.close()


@pytest.mark.parametrize("ni", ["0", "2", "5"])
def test_imad_max_fast_kernel_gives_identical_candidates(ob, monkeypatch, ni):
    """The default k_fast_t<1, 8, 8> takes 16 of the 72 min/max per pixel pair from max(a, b) = a + b - min(a, b) on the
    packed words (two IMADs on the FMA pipe); ORBX_FAST1_NI=0 / 2 / 5 select the (0, 0) / (8, 0) / (8, 4) instantiations.
    All must give the same bytes as the default and the oracle."""
    monkeypatch.setenv("ORBX_FAST1_NI", ni)
    p = _params(1000)
    exv = ob.ORBextractor(**p)
    monkeypatch.delenv("ORBX_FAST1_NI")
    ex1 = ob.ORBextractor(**p)
    orc = Oracle(**p)
    for (w, h), seed, contrast in (((640, 480), 7, 1.0), ((640, 480), 8, 0.2), ((757, 483), 9, 1.0), ((480, 640), 10, 1.0)):
        img = synth_image(w, h, seed).astype(np.float32)
        img = np.clip((img - 128.0) * contrast + 128.0, 0, 255).astype(np.uint8)
        kv, dv = exv(img); k1, d1 = ex1(img)
        assert k1.tobytes() == kv.tobytes() and np.array_equal(d1, dv)
        _compare_frame(exv, orc, img, 0, None, kv, dv)
    ex1.close(); exv.close()


def test_single_process_multi_gpu_entry_point(ob):
    """orbx_extract_batch_multi: contiguous blocks of the batch on several handles (every visible GPU, at least two handles
    so that a one-GPU box still exercises the split), results in disjoint slices — bytes equal to one device's call."""
    import torch
    from orbslam2_b200.sharding import MultiGpuExtractor, shard_sizes
    p = _params(1000)
    imgs = np.stack([synth_image(640, 480, 40 + s) for s in range(7)])
    ex = ob.ORBextractor(**p)
    k1, d1, n1 = ex.extract_batch(imgs)
    ndev = torch.cuda.device_count()
    devices = list(range(ndev)) if ndev >= 2 else [0, 0, 0]
    assert shard_sizes(7, len(devices))[0] >= 1
    mg = MultiGpuExtractor(devices, **p)
    k2, d2, n2 = mg.extract_batch(imgs)
    assert np.array_equal(n1, n2)
    for f in range(7):
        assert k1[f, :n1[f]].tobytes() == k2[f, :n2[f]].tobytes() and np.array_equal(d1[f, :n1[f]], d2[f, :n2[f]])
    # fewer frames than handles: the empty blocks are skipped
    k3, d3, n3 = mg.extract_batch(imgs[:2])
    assert np.array_equal(n3, n1[:2]) and k3[1, :n3[1]].tobytes() == k1[1, :n1[1]].tobytes()
    mg.close(); ex.close()
