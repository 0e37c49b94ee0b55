"""CPU: the vocabulary-tree oracle (oracle/orb_oracle.cpp orc_fbow_*) pinned against the reference's own fbow code —
golden outputs of the unmodified fbow.cpp (tests/golden/fbow_ref.npz, made by tests/golden/make_fbow_golden.py) and,
where oracle/_ref/libref_fbow.so is present, the library itself on further cases. Plus the host-side error behaviour of
orbv_create (no GPU needed: the blob is validated before CUDA is touched)."""
import ctypes as C
import hashlib
import os
import sys

import numpy as np
import pytest

from oracle_lib import OracleVocabulary, RefVocabulary, have_ref_fbow
from orbslam2_b200.synth import synth_vocabulary, synth_vocabulary_features

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_fbow_golden import CASES, LEVELS, N_FEATURES  # noqa: E402


def flat(res):
    (wid, ww), fv = res
    keys = np.array(sorted(fv), np.uint32)
    noff = np.cumsum([0] + [len(fv[int(k)]) for k in keys]).astype(np.int32)
    feat = np.concatenate([fv[int(k)] for k in keys]).astype(np.uint32) if len(keys) else np.zeros(0, np.uint32)
    return wid, ww, keys, noff, feat


def same(a, b):
    return all(x.shape == y.shape and x.tobytes() == y.tobytes() for x, y in zip(flat(a), flat(b)))


def test_oracle_matches_reference_golden():
    G = np.load(os.path.join(HERE, "golden", "fbow_ref.npz"))
    for c, kw in enumerate(CASES):
        blob = synth_vocabulary(**kw)
        assert hashlib.sha256(blob).digest() == G[f"c{c}_sha"].tobytes(), "synthetic vocabulary generator changed"
        orc = OracleVocabulary(blob)
        feats = synth_vocabulary_features(blob, N_FEATURES, seed=10 + c)
        for lvl in LEVELS:
            got = flat(orc.transform(feats, lvl))
            for name, arr in zip(("wid", "ww", "nid", "noff", "feat"), got):
                exp = G[f"c{c}_l{lvl}_{name}"]
                assert arr.shape == exp.shape and arr.tobytes() == exp.tobytes(), (c, lvl, name)
        wid, ww = orc.transform(feats)
        assert np.array_equal(wid, G[f"c{c}_norm_wid"]) and ww.tobytes() == G[f"c{c}_norm_ww"].tobytes()
        half = orc.transform(feats[:N_FEATURES // 2])
        assert np.array_equal(np.array([orc.score((wid, ww), half), orc.score(half, half)]), G[f"c{c}_score"])


@pytest.mark.skipif(not have_ref_fbow(), reason="oracle/_ref/libref_fbow.so not built (needs /root/reference)")
@pytest.mark.parametrize("kw", [dict(k=10, levels=6, ragged=True, leaf_prob=0.2, align=32, seed=5),
                                dict(k=16, levels=2, align=4, seed=6), dict(k=2, levels=9, seed=7),
                                dict(k=13, levels=3, ragged=True, seed=8)])
def test_oracle_matches_reference_library(kw):
    blob = synth_vocabulary(**kw)
    orc, ref = OracleVocabulary(blob), RefVocabulary(blob)
    rng = np.random.default_rng(kw["seed"])
    feats = np.concatenate([synth_vocabulary_features(blob, 1500, seed=kw["seed"]),
                            rng.integers(0, 256, (500, 32), dtype=np.uint8)])
    for lvl in (0, 1, 3, 4, 12):
        assert same(orc.transform(feats, lvl), ref.transform(feats, lvl)), lvl
    a, b = orc.transform(feats), ref.transform(feats)
    assert np.array_equal(a[0], b[0]) and a[1].tobytes() == b[1].tobytes()
    c = orc.transform(feats[::3])
    assert orc.score(a, c) == ref.score(a, c) and orc.score(c, c) == ref.score(c, c)
    # single feature; duplicated features (weights of one word added twice: a float sum, not a product)
    assert same(orc.transform(feats[:1], 4), ref.transform(feats[:1], 4))
    dup = np.repeat(feats[:7], 5, axis=0)
    assert same(orc.transform(dup, 4), ref.transform(dup, 4))


def test_vocabulary_blob_validation_needs_no_gpu():
    """orbv_create rejects malformed blobs with ORBX_EINVAL before touching CUDA, and on a box without a GPU a valid
    blob fails loudly with ORBX_ECUDA (there is no CPU fallback)."""
    import torch
    from orbslam2_b200 import _lib
    L = _lib.lib()
    blob = synth_vocabulary(k=10, levels=2, seed=0)
    h = C.c_void_p()

    def create(b):
        return L.orbv_create(bytes(b), len(b), 0, C.byref(h))

    assert create(blob[:64]) == _lib.ORBX_EINVAL
    bad = bytearray(blob); bad[0] ^= 1
    assert create(bad) == _lib.ORBX_EINVAL and b"signature" in L.orbx_last_error_string()
    bad = bytearray(blob); bad[8 + 108] = 61            # AKAZE-sized descriptors
    assert create(bad) == _lib.ORBX_EINVAL
    bad = bytearray(blob); bad[8 + 56] += 1             # one block more than the data holds
    assert create(bad) == _lib.ORBX_EINVAL
    assert create(blob[:-1]) == _lib.ORBX_EINVAL         # truncated
    bad = bytearray(blob); bad[128] = 11                # block 0 claims 11 children with k = 10
    assert create(bad) == _lib.ORBX_EINVAL
    if not torch.cuda.is_available():
        assert create(blob) == _lib.ORBX_ECUDA


# ---- oracle of ORBmatcher::SearchByFboW on hand-made cases with known answers ------------------------------------------
def _side(desc, fv, elig=None, angle=None):
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    keys = sorted(fv)
    off = np.zeros(len(keys) + 1, np.int32)
    off[1:] = np.cumsum([len(fv[k]) for k in keys])
    feat = np.concatenate([np.asarray(fv[k], np.uint32) for k in keys]) if keys else np.zeros(0, np.uint32)
    return dict(n=len(desc), desc=desc, angle=np.zeros(len(desc), np.float32) if angle is None else np.asarray(angle, np.float32),
                eligible=None if elig is None else np.asarray(elig, np.uint8), node_id=np.array(keys, np.uint32),
                node_off=off, node_feat=np.ascontiguousarray(feat, np.uint32))


def _with_bits(n):
    d = np.zeros(32, np.uint8)
    d[:n // 8] = 255
    if n % 8:
        d[n // 8] = (1 << (n % 8)) - 1
    return d


def test_search_by_bow_oracle_known_answers():
    from oracle_lib import oracle_search_by_bow
    z = _with_bits(0)
    # B: distances 0, 10, 40 from the zero descriptor, one far away (200)
    B = _side([_with_bits(0), _with_bits(10), _with_bits(40), _with_bits(200)], {3: [0, 1, 2], 5: [3]})
    # A: two zero descriptors in node 3 (the 2nd finds B0 taken and falls back to B1: 10 < 0.7*40), one in node 4 (no
    # such node in B), one ineligible
    A = _side([z, z, z, z], {3: [0, 1, 3], 4: [2]}, elig=[1, 1, 1, 0])
    n, m = oracle_search_by_bow(A, B, 0.7, 50, False, False)
    assert n == 2 and m.tolist() == [0, 1, -1, -1]
    # ratio test: best 10, second 12 -> rejected at 0.7, accepted at 0.9
    B2 = _side([_with_bits(10), _with_bits(12)], {3: [0, 1]})
    A2 = _side([z], {3: [0]}, elig=[1])
    assert oracle_search_by_bow(A2, B2, 0.7, 50, False, False)[0] == 0
    assert oracle_search_by_bow(A2, B2, 0.9, 50, False, False)[1].tolist() == [0]
    # threshold: best == 50 passes `<=` (KF, F) and fails `<` (KF, KF)
    B3 = _side([_with_bits(50)], {3: [0]})
    assert oracle_search_by_bow(A2, B3, 0.9, 50, False, False)[0] == 1
    assert oracle_search_by_bow(A2, _side([_with_bits(50)], {3: [0]}, elig=[1]), 0.9, 50, True, False)[0] == 0
    # KF, KF: a candidate without a map point is invisible (neither best nor second)
    B4 = _side([_with_bits(2), _with_bits(3)], {3: [0, 1]}, elig=[0, 1])
    assert oracle_search_by_bow(A2, B4, 0.7, 50, True, False)[1].tolist() == [1]
    # rotation filter: 12 matches at a 30 degree offset, one at 200 -> the lone bin is dropped (< 10 % of the peak)
    nA = 13
    dA = np.stack([_with_bits(8 * i) for i in range(nA)])          # mutually far apart (>= 8 bits)
    ang_a = np.full(nA, 100, np.float32)
    ang_b = np.full(nA, 70, np.float32); ang_b[12] = 260
    fv = {i: [i] for i in range(nA)}
    n, m = oracle_search_by_bow(_side(dA, fv, elig=np.ones(nA), angle=ang_a), _side(dA, fv, angle=ang_b), 0.9, 50, False, True)
    assert n == 12 and m[12] == -1 and m[:12].tolist() == list(range(12))
