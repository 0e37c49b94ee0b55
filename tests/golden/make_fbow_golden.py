#!/usr/bin/env python
"""Generates tests/golden/fbow_ref.npz: outputs of the UNMODIFIED reference fbow::Vocabulary
(/root/reference/Thirdparty/fbow/src/fbow.cpp compiled by oracle/Makefile into oracle/_ref/libref_fbow.so) on
deterministic synthetic vocabularies and descriptors. Run in the build container (needs /root/reference):

    python tests/golden/make_fbow_golden.py

The vocabularies and descriptors are NOT stored: tests regenerate them from the same seeds (orbslam2_b200/synth.py) and
the fixture holds a digest of the blob so that a change of the generator is noticed."""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle_lib import RefVocabulary  # noqa: E402
from orbslam2_b200.synth import synth_vocabulary, synth_vocabulary_features  # noqa: E402

CASES = [dict(k=10, levels=3, seed=1), dict(k=10, levels=5, seed=2, ragged=True, leaf_prob=0.15, align=32),
         dict(k=6, levels=4, seed=3, ragged=True, leaf_prob=0.1, align=16)]
N_FEATURES, LEVELS = 700, (0, 2, 4)


def main():
    out = {}
    for c, kw in enumerate(CASES):
        blob = synth_vocabulary(**kw)
        out[f"c{c}_sha"] = np.frombuffer(hashlib.sha256(blob).digest(), np.uint8)
        ref = RefVocabulary(blob)
        feats = synth_vocabulary_features(blob, N_FEATURES, seed=10 + c)
        for lvl in LEVELS:
            (wid, ww), fv = ref.transform(feats, lvl)
            keys = np.array(sorted(fv), np.uint32)
            out[f"c{c}_l{lvl}_wid"], out[f"c{c}_l{lvl}_ww"] = wid, ww
            out[f"c{c}_l{lvl}_nid"] = keys
            out[f"c{c}_l{lvl}_noff"] = np.cumsum([0] + [len(fv[int(k)]) for k in keys]).astype(np.int32)
            out[f"c{c}_l{lvl}_feat"] = np.concatenate([fv[int(k)] for k in keys]).astype(np.uint32)
        wid, ww = ref.transform(feats)
        out[f"c{c}_norm_wid"], out[f"c{c}_norm_ww"] = wid, ww
        half = ref.transform(feats[:N_FEATURES // 2])
        out[f"c{c}_score"] = np.array([ref.score((wid, ww), half), ref.score(half, half)], np.float64)
    np.savez_compressed(os.path.join(HERE, "fbow_ref.npz"), **out)
    print("wrote fbow_ref.npz:", len(out), "arrays")


if __name__ == "__main__":
    main()
