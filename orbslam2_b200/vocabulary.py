"""Host-side mirror of fbow::Vocabulary / fbow::fBow / fbow::fBow2 as the reference uses them
(/root/reference/Thirdparty/fbow/src/fbow.h:31-125, fbow.cpp; callers src/Frame.cc:395-400, src/KeyFrame.cc:62,
src/KeyFrameDatabase.cc:130,247). All compute runs in liborbx's orbv_* kernels; there is no CPU path.

fBow  (std::map<uint32_t, float>)              -> (ids uint32 ascending, weights float32)
fBow2 (std::map<uint32_t, vector<uint32_t>>)  -> {node id: feature indices uint32 ascending}, keys ascending
"""
import ctypes as C

import numpy as np

from . import _lib


def _is_torch(x):
    return type(x).__module__.startswith("torch")


class Vocabulary:
    def __init__(self, device=0):
        self._L = _lib.lib()
        self._h = None
        self.device = device
        self._spaces = (_lib.MEM_HOST, _lib.MEM_HOST)

    # --- loading (fbow.cpp:156-183) ----------------------------------------------------------------------------------
    def readFromFile(self, filepath):
        try:
            with open(filepath, "rb") as f:
                blob = f.read()
        except OSError:
            raise RuntimeError("Vocabulary::readFromFile could not open:" + str(filepath))
        self.fromStream(blob)

    def fromStream(self, blob):
        self.clear()
        h = C.c_void_p()
        blob = bytes(blob)
        _lib.check(self._L.orbv_create(blob, len(blob), self.device, C.byref(h)))
        self._h = h
        self._spaces = (_lib.MEM_HOST, _lib.MEM_HOST)

    def clear(self):
        if self._h is not None:
            self._L.orbv_destroy(self._h)
            self._h = None

    close = clear

    def __del__(self):
        try:
            self.clear()
        except Exception:
            pass

    # --- getters (fbow.h:96-108) ---------------------------------------------------------------------------------------
    def isValid(self):
        return self._h is not None

    def _info(self):
        self._need()
        k, nb, ds, tw = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint64()
        _lib.check(self._L.orbv_info(self._h, C.byref(k), C.byref(nb), C.byref(ds), C.byref(tw)))
        return k.value, nb.value, ds.value, tw.value

    def getK(self):
        return self._info()[0]

    def size(self):
        return self._info()[1]

    def getDescSize(self):
        return self._info()[2]

    def getTotalWords(self):
        return self._info()[3]

    def _need(self):
        if self._h is None:
            raise RuntimeError("Vocabulary is not loaded")

    def _set_space(self, i, o):
        if (i, o) != self._spaces:
            _lib.check(self._L.orbv_set_memory_space(self._h, i, o))
            self._spaces = (i, o)

    # --- transform -----------------------------------------------------------------------------------------------------
    def transform(self, features, level=None):
        """level given: transform(features, level, fBow&, fBow2&) -> ((ids, weights), {node: features}) (fbow.cpp:52-93);
        level None: transform(features) -> L2-normalised (ids, weights) (fbow.cpp:95-153)."""
        features = np.ascontiguousarray(features, np.uint8)
        if features.size == 0:
            raise RuntimeError("Vocabulary::transform No input data")
        if features.ndim != 2 or features.shape[1] != 32:
            raise RuntimeError("Vocabulary::transform features are of different size than the vocabulary ones")
        n = np.array([len(features)], np.int32)
        res = self.transform_batch(features[None], n, level)
        return res[0]

    def transform_batch(self, desc, n, level=None):
        """desc [B,cap,32] uint8 (numpy, or a torch CUDA tensor such as ORBextractor.extract_batch returns), n [B] rows
        per frame. Returns one result per frame, as transform() does. With a CUDA tensor the descriptors never leave
        the device; the (small) results always come back to the host."""
        self._need()
        n = np.ascontiguousarray(n, np.int32)
        B, cap = int(desc.shape[0]), int(desc.shape[1])
        assert desc.shape[2] == 32 and len(n) == B
        if _is_torch(desc):
            import torch
            assert desc.is_cuda and desc.dtype == torch.uint8 and desc.is_contiguous() and desc.device.index == self.device
            torch.cuda.current_stream(desc.device).synchronize()
            self._set_space(_lib.MEM_DEVICE, _lib.MEM_HOST)
            dptr = desc.data_ptr()
        else:
            desc = np.ascontiguousarray(desc, np.uint8)
            self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
            dptr = desc.ctypes.data
        wid, ww = np.zeros((B, cap), np.uint32), np.zeros((B, cap), np.float32)
        nid, noff, feat = np.zeros((B, cap), np.uint32), np.zeros((B, cap + 1), np.int32), np.zeros((B, cap), np.uint32)
        nw, nn = np.zeros(B, np.int32), np.zeros(B, np.int32)
        lvl = -1 if level is None else int(level)
        _lib.check(self._L.orbv_transform_batch(self._h, dptr, n.ctypes.data, B, cap, lvl, int(level is None),
                                                wid.ctypes.data, ww.ctypes.data, nw.ctypes.data, nid.ctypes.data,
                                                noff.ctypes.data, feat.ctypes.data, nn.ctypes.data))
        out = []
        for f in range(B):
            bow = (wid[f, :nw[f]].copy(), ww[f, :nw[f]].copy())
            if level is None:
                out.append(bow)
            else:
                out.append((bow, {int(nid[f, j]): feat[f, noff[f, j]:noff[f, j + 1]].copy() for j in range(nn[f])}))
        return out

    # --- fBow::score (fbow.cpp:208-254) ----------------------------------------------------------------------------------
    def score_batch(self, q, db):
        """score(q, v) for every v in db (a list of (ids, weights)); returns float64 [len(db)]."""
        self._need()
        m = len(db)
        off = np.zeros(m + 1, np.int32)
        off[1:] = np.cumsum([len(v[0]) for v in db])
        did = np.ascontiguousarray(np.concatenate([v[0] for v in db]) if m else np.zeros(0), np.uint32)
        dw = np.ascontiguousarray(np.concatenate([v[1] for v in db]) if m else np.zeros(0), np.float32)
        qi, qw = np.ascontiguousarray(q[0], np.uint32), np.ascontiguousarray(q[1], np.float32)
        s = np.zeros(m, np.float64)
        _lib.check(self._L.orbv_score_batch(self._h, qi.ctypes.data, qw.ctypes.data, len(qi), did.ctypes.data,
                                            dw.ctypes.data, off.ctypes.data, m, s.ctypes.data))
        return s

    def score(self, a, b):
        return float(self.score_batch(a, [b])[0])

    def launch_count(self):
        return int(self._L.orbv_launch_count(self._h)) if self._h is not None else 0

    def cuda_stream(self):
        s = C.c_void_p()
        _lib.check(self._L.orbv_stream(self._h, C.byref(s)))
        return s.value
