"""Multi-GPU sharding of the hot path (SURVEY.md §8 e): frames (and query-descriptor blocks) are independent units, so
they are split into contiguous blocks, one block per rank / GPU, with NO collective on the data path. The only exchange
is the optional host-side gather of the per-rank result rows, done with torch.distributed on whatever backend the
caller initialised (NCCL on the B200 box, gloo in the CPU tests).

The reference has no counterpart (single process, std::thread; /root/reference/src/System.cc:104-112): a stereo pair is
kept on ONE rank so Frame::ComputeStereoMatches (/root/reference/src/Frame.cc:464-642) reads both pyramids locally.
"""
import numpy as np


def shard_range(n_units, rank, world):
    """Contiguous block [lo, hi) of `n_units` owned by `rank`; the first n_units % world ranks get one extra unit."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, extra = divmod(n_units, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_sizes(n_units, world):
    return [shard_range(n_units, r, world)[1] - shard_range(n_units, r, world)[0] for r in range(world)]


def extract_sharded(extract_fn, images, rank, world):
    """Run `extract_fn(images[lo:hi]) -> (kps[b,cap], desc[b,cap,32], n[b])` on this rank's block of frames."""
    lo, hi = shard_range(len(images), rank, world)
    if hi == lo:
        return lo, hi, None
    return lo, hi, extract_fn(images[lo:hi])


def gather_frame_results(local, n_frames, cap, kp_dtype, group=None):
    """Host-side concatenation of per-rank results into full arrays on every rank (all_gather_object: the result rows
    are small — 60 B per keypoint). `local` is the third element returned by extract_sharded (or None)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    parts = [None] * world
    payload = None
    if local is not None:
        k, d, n = local
        payload = (np.asarray(k), np.asarray(d), np.asarray(n))
    dist.all_gather_object(parts, payload, group=group)
    kps = np.zeros((n_frames, cap), kp_dtype)
    desc = np.zeros((n_frames, cap, 32), np.uint8)
    cnt = np.zeros(n_frames, np.int32)
    for r, part in enumerate(parts):
        lo, hi = shard_range(n_frames, r, world)
        if part is None:
            continue
        k, d, n = part
        c = min(cap, k.shape[1])
        kps[lo:hi, :c] = k[:, :c]
        desc[lo:hi, :c] = d[:, :c]
        cnt[lo:hi] = n
    return kps, desc, cnt


def knn2_sharded(knn2_fn, queries, rank, world):
    """Config-4 sharding: the train set is replicated, queries are split; returns (lo, hi, (i1, d1, d2))."""
    lo, hi = shard_range(len(queries), rank, world)
    if hi == lo:
        return lo, hi, None
    return lo, hi, knn2_fn(queries[lo:hi])


class MultiGpuExtractor:
    """Single-process multi-GPU extraction (orbx_extract_batch_multi, include/orbx.h): one ORBextractor handle per device,
    each driven by its own host thread inside the C library; frames are split into the contiguous shard_range blocks and
    the results land in disjoint slices of ONE set of host arrays. `devices` may name a device more than once (two
    handles on one GPU: what a single-GPU box can test)."""

    def __init__(self, devices, **params):
        from .extractor import ORBextractor
        self.ex = [ORBextractor(device=d, **params) for d in devices]

    def extract_batch(self, images):
        import ctypes as C
        from . import _lib
        from .extractor import KP_DTYPE
        images = np.ascontiguousarray(images, np.uint8)
        B, H, W = images.shape
        cap = self.ex[0].max_keypoints(W, H)
        kps = np.zeros((B, cap), KP_DTYPE); desc = np.zeros((B, cap, 32), np.uint8); n = np.zeros(B, np.int32)
        for e in self.ex:
            e._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        hs = (C.c_void_p * len(self.ex))(*[e._h for e in self.ex])
        _lib.check(self.ex[0]._L.orbx_extract_batch_multi(hs, len(self.ex), images.ctypes.data, B, W, H, W, W * H,
                                                          kps.ctypes.data, desc.ctypes.data, cap, n.ctypes.data))
        return kps, desc, n

    def close(self):
        for e in self.ex:
            e.close()
