"""GPU parity: the device vocabulary-tree transform (orbv_* entry points through orbslam2_b200.Vocabulary) vs the CPU
oracle (orc_fbow_*, itself pinned against the unmodified reference fbow.cpp in tests/test_cpu_fbow.py).

Bar: bit-exact — word ids, node ids, feature lists, and the float32 weight sums / float64 scores bit for bit."""
import numpy as np
import pytest

from oracle_lib import OracleVocabulary, TUM1
from orbslam2_b200.synth import synth_image, synth_vocabulary, synth_vocabulary_features
from test_cpu_fbow import same

pytestmark = pytest.mark.gpu

VOCABS = [dict(k=10, levels=4, seed=1), dict(k=10, levels=6, seed=2, ragged=True, leaf_prob=0.15, align=32),
          dict(k=6, levels=4, seed=3, ragged=True, leaf_prob=0.1, align=16), dict(k=16, levels=2, seed=4, align=4),
          dict(k=2, levels=9, seed=5)]


@pytest.fixture(scope="module")
def ob():
    import orbslam2_b200
    return orbslam2_b200


def _features(blob, n, seed):
    rng = np.random.default_rng(seed)
    f = synth_vocabulary_features(blob, n, seed=seed)
    f[rng.random(n) < 0.05] = f[0]      # repeated descriptors: one word hit several times
    return f


@pytest.mark.parametrize("kw", VOCABS)
def test_transform_levels_vs_oracle(ob, kw):
    blob = synth_vocabulary(**kw)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    orc = OracleVocabulary(blob)
    feats = _features(blob, 1800, kw["seed"])
    for lvl in (0, 1, 2, 4, 12):
        assert same(voc.transform(feats, lvl), orc.transform(feats, lvl)), lvl
    a, b = voc.transform(feats), orc.transform(feats)
    assert np.array_equal(a[0], b[0]) and a[1].tobytes() == b[1].tobytes(), "normalised fBow"
    voc.close()


def test_ragged_batch_and_sort_sizes(ob):
    """One call over frames with 0, 1, 2, 3, 1023..1025 and 5000 features (every bitonic-sort size incl. the padding
    cases); results equal the oracle frame by frame and the single-frame call."""
    blob = synth_vocabulary(k=10, levels=5, seed=9, ragged=True, leaf_prob=0.05)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    orc = OracleVocabulary(blob)
    counts = [0, 1, 2, 3, 700, 1023, 1024, 1025, 5000, 17]
    cap = max(counts)
    desc = np.zeros((len(counts), cap, 32), np.uint8)
    for f, n in enumerate(counts):
        if n:
            desc[f, :n] = _features(blob, n, 100 + f)
    res = voc.transform_batch(desc, counts, 4)
    for f, n in enumerate(counts):
        if n == 0:
            assert len(res[f][0][0]) == 0 and res[f][1] == {}
            continue
        assert same(res[f], orc.transform(desc[f, :n], 4)), f"frame {f} ({n} features)"
    assert same(voc.transform(desc[4, :700], 4), res[4])
    with pytest.raises(RuntimeError):
        voc.transform(np.zeros((0, 32), np.uint8), 4)       # "No input data" (fbow.cpp:53)
    with pytest.raises(ob.OrbxError):
        voc.transform_batch(np.zeros((1, 20000, 32), np.uint8), [20000], 4)   # above the per-frame limit
    voc.close()


def test_getters_and_score(ob):
    blob = synth_vocabulary(k=10, levels=3, seed=11)
    voc = ob.Vocabulary(); assert not voc.isValid()
    voc.fromStream(blob)
    assert voc.isValid() and voc.getK() == 10 and voc.size() == 111 and voc.getDescSize() == 32
    assert voc.getTotalWords() == 1000
    orc = OracleVocabulary(blob)
    vecs = [voc.transform(_features(blob, n, 200 + i)) for i, n in enumerate((900, 1200, 40, 1, 1500))]
    q = vecs[0]
    got = voc.score_batch(q, vecs)
    exp = np.array([orc.score(q, v) for v in vecs])
    assert got.tobytes() == exp.tobytes()
    assert got[0] == 1.0 or abs(got[0] - 1.0) < 1e-6
    assert voc.score(vecs[1], vecs[2]) == orc.score(vecs[1], vecs[2])
    empty = (np.zeros(0, np.uint32), np.zeros(0, np.float32))
    assert voc.score(q, empty) == orc.score(q, empty) == 0.0
    voc.close(); assert not voc.isValid()


def test_device_descriptors_from_extractor(ob):
    """Frame::ComputeFboW on the extractor's device output: the descriptors go from orbx_extract_batch to
    orbv_transform_batch without leaving the GPU and give the result of the host path."""
    import torch
    blob = synth_vocabulary(k=10, levels=4, seed=12)
    voc = ob.Vocabulary(); voc.fromStream(blob)
    orc = OracleVocabulary(blob)
    ex = ob.ORBextractor(**TUM1)
    imgs = torch.from_numpy(np.stack([synth_image(640, 480, s) for s in range(3)])).cuda()
    kps, desc, n = ex.extract_batch(imgs)
    res = voc.transform_batch(desc, n, 4)
    dh = desc.cpu().numpy()
    for f in range(3):
        assert same(res[f], orc.transform(dh[f, :n[f]], 4))
    assert same(voc.transform(dh[1, :n[1]], 4), res[1])
    assert voc.launch_count() > 0
    ex.close(); voc.close()
