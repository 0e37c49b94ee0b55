"""Developer diagnostic (GPU box): run every stage on a few inputs and print ALL mismatches against the oracle."""
import sys, os, time, traceback
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from oracle_lib import Oracle, OraclePrims, canonical, TUM1
from orbslam2_b200.synth import synth_image, synth_descriptors, synth_stereo_pair
import orbslam2_b200 as ob


def check_extract(w, h, nf, seeds=(0,)):
    p = dict(TUM1); p["nfeatures"] = nf
    ex = ob.ORBextractor(**p); orc = Oracle(**p)
    for seed in seeds:
        img = synth_image(w, h, seed)
        t = time.time(); kps, desc = ex(img); dt = time.time() - t
        ko, do = orc.extract(img, 4 * nf); ko, do = canonical(ko, do)
        print(f"[{w}x{h} seed {seed}] gpu n={len(kps)} oracle n={len(ko)} ({dt*1e3:.1f} ms first call)")
        for l in range(8):
            a = ex.pyramid_level(l, 0, True); b = orc.level(l, True)
            bl = ex.blurred_level(l, 0); bo = orc.blurred(l)
            cg = ex.candidates(l, 0); co = orc.candidates(l)
            cg = cg[np.lexsort((cg[:, 0], cg[:, 1]))]; co = co[np.lexsort((co[:, 0], co[:, 1]))]
            sg = set(map(tuple, cg.tolist())); so = set(map(tuple, co.tolist()))
            kg = kps[kps["octave"] == l]; kk = ko[ko["octave"] == l]
            print(f"  L{l} {a.shape}: pyr diff {(a != b).sum()}  blur diff {(bl != bo).sum()}  cand gpu {len(cg)} orc {len(co)} "
                  f"common {len(sg & so)}  sel gpu {len(kg)} orc {len(kk)}")
        if len(kps) == len(ko):
            print("  coords max diff", np.abs(kps["x"] - ko["x"]).max(), np.abs(kps["y"] - ko["y"]).max(),
                  "angle max diff", np.abs(kps["angle"] - ko["angle"]).max(),
                  "resp eq", np.array_equal(kps["response"], ko["response"]), "size eq", np.array_equal(kps["size"], ko["size"]),
                  "desc identical", int((desc == do).all(axis=1).sum()), "/", len(do))
        else:
            sg = set(zip(kps["octave"].tolist(), kps["x"].tolist(), kps["y"].tolist()))
            so = set(zip(ko["octave"].tolist(), ko["x"].tolist(), ko["y"].tolist()))
            print("  selected sets: common", len(sg & so), "gpu-only", len(sg - so), "oracle-only", len(so - sg))
    ex.set_profiling(True)
    imgs = np.stack([synth_image(w, h, s) for s in range(8)])
    for _ in range(2):
        ex.extract_batch(imgs)
    print("  stage ms (batch 8):", ex.stage_ms())
    ex.close()


def check_match():
    m = ob.ORBmatcher(); P = OraclePrims()
    t, q = synth_descriptors(20000, 1500, seed=1, dup=100)
    i1, d1, d2 = m.knn2(q, t); o1, e1, e2 = P.knn2(q, t, popcnt=True)
    print("knn2 mismatches", (i1 != o1).sum(), (d1 != e1).sum(), (d2 != e2).sum())
    t, q = synth_descriptors(1000000, 4096, seed=2)
    m.set_train(t)
    for _ in range(2):
        t0 = time.time(); m.knn2(q); dt = time.time() - t0
    print(f"knn2 4096 x 1M: {dt*1e3:.1f} ms incl. copies -> {4096*1e6/dt/1e9:.1f} Gpairs/s")


def check_stereo():
    p = dict(TUM1); p["nfeatures"] = 2000
    exL = ob.ORBextractor(**p); exR = ob.ORBextractor(**p); P = OraclePrims()
    a, b = synth_stereo_pair(1241, 376, 0)
    kL, dL, nL = exL.extract_batch(a[None]); kR, dR, nR = exR.extract_batch(b[None])
    u, z = ob.compute_stereo_matches(exL, exR, 386.1448, 0.0, 718.856)
    pyrL = [exL.pyramid_level(l, 0) for l in range(8)]; pyrR = [exR.pyramid_level(l, 0) for l in range(8)]
    eu, ez, sad, n = P.stereo_match(kL[0, :nL[0]], dL[0, :nL[0]], kR[0, :nR[0]], dR[0, :nR[0]], pyrL, pyrR,
                                    exL.GetScaleFactors(), exL.GetInverseScaleFactors(), 386.1448, 0.0, 718.856)
    print("stereo: matches oracle", int((eu >= 0).sum()), "gpu", int((u[0, :nL[0]] >= 0).sum()),
          "uRight diff", int((u[0, :nL[0]] != eu).sum()), "depth diff", int((z[0, :nL[0]] != ez).sum()))


for fn, args in ((check_extract, (640, 480, 1000, (0, 1))), (check_extract, (1241, 376, 2000)), (check_match, ()), (check_stereo, ())):
    try:
        fn(*args)
    except Exception:
        traceback.print_exc()
