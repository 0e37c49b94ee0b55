"""ctypes bindings for the CPU checkers under oracle/ (test infrastructure only).

`Oracle`  -> oracle/liborb_oracle.so          (our restatement, built by oracle/Makefile)
`RefExtractor` -> oracle/_ref/libref_orbextractor.so (the unmodified reference ORBextractor.cc on the cv shim)
"""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liborb_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_orbextractor.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

TUM1 = dict(nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7, patchSize=31, halfPatchSize=15,
            edgeThreshold=19)

_u8p = C.POINTER(C.c_uint8)


def _p(a, t=C.c_void_p):
    return a.ctypes.data_as(t)


def build_oracle():
    if not os.path.exists(ORACLE_SO) or (os.path.isdir("/root/reference") and not os.path.exists(REF_SO)):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL)


def _load(path):
    lib = C.CDLL(path)
    return lib


class _ExtractorBase:
    prefix = None

    def __init__(self, lib, **kw):
        p = dict(TUM1); p.update(kw)
        self.params = p
        self.lib = lib
        f = getattr(lib, self.prefix + "_create"); f.restype = C.c_void_p
        f.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        self.h = C.c_void_p(f(p["nfeatures"], p["scaleFactor"], p["nlevels"], p["iniThFAST"], p["minThFAST"],
                              p["patchSize"], p["halfPatchSize"], p["edgeThreshold"]))
        self._fn("destroy").argtypes = [C.c_void_p]
        e = self._fn("extract"); e.restype = C.c_int
        e.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int]
        self._fn("level_size").argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        self._fn("level_copy").argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        self._fn("tables").argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]

    def _fn(self, name):
        return getattr(self.lib, self.prefix + "_" + name)

    def __del__(self):
        try:
            self._fn("destroy")(self.h)
        except Exception:
            pass

    def extract(self, img, cap=None):
        img = np.ascontiguousarray(img, np.uint8)
        h, w = img.shape
        cap = cap or (self.params["nfeatures"] * 2 + 64)
        kps = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
        n = self._fn("extract")(self.h, _p(img), w, h, w, _p(kps), _p(desc), cap)
        assert n >= 0, "capacity too small"
        return kps[:n].copy(), desc[:n].copy()

    def level_size(self, l):
        w = C.c_int(); h = C.c_int()
        self._fn("level_size")(self.h, l, C.byref(w), C.byref(h))
        return w.value, h.value

    def level(self, l, with_border=False):
        w, h = self.level_size(l)
        e = self.params["edgeThreshold"] if with_border else 0
        out = np.zeros((h + 2 * e, w + 2 * e), np.uint8)
        self._fn("level_copy")(self.h, l, _p(out), int(with_border))
        return out

    def tables(self):
        n = self.params["nlevels"]
        t = np.zeros(4 * n, np.float32); pl = np.zeros(n, np.int32); um = np.zeros(self.params["halfPatchSize"] + 1, np.int32)
        self._fn("tables")(self.h, _p(t), _p(pl), _p(um))
        return t.reshape(4, n), pl, um


class Oracle(_ExtractorBase):
    prefix = "orc"

    def __init__(self, **kw):
        build_oracle()
        lib = _load(ORACLE_SO)
        super().__init__(lib, **kw)
        lib.orc_cand_count.argtypes = [C.c_void_p, C.c_int]
        lib.orc_cand_copy.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        lib.orc_sel_count.argtypes = [C.c_void_p, C.c_int]
        lib.orc_blur_copy.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        lib.orc_set_tiebreak.argtypes = [C.c_void_p, C.c_int]

    def candidates(self, l):
        n = self.lib.orc_cand_count(self.h, l)
        out = np.zeros((n, 3), np.int32)
        if n:
            self.lib.orc_cand_copy(self.h, l, _p(out))
        return out

    def blurred(self, l):
        w, h = self.level_size(l)
        out = np.zeros((h, w), np.uint8)
        self.lib.orc_blur_copy(self.h, l, _p(out))
        return out

    def tie_at_cut(self, l):
        """True when the last sorted quadtree sweep of level l was cut between two nodes of equal key count: the only
        case in which the reference's selection depends on heap addresses (src/ORBextractor.cc:678)."""
        return bool(self.lib.orc_tie_at_cut(self.h, l))

    def set_tiebreak(self, m):
        self.lib.orc_set_tiebreak(self.h, m)


class RefExtractor(_ExtractorBase):
    prefix = "ref"

    def __init__(self, **kw):
        build_oracle()
        super().__init__(_load(REF_SO), **kw)


def have_ref():
    build_oracle()
    return os.path.exists(REF_SO)


class OraclePrims:
    """Stand-alone primitives + matcher loops of the oracle."""

    def __init__(self):
        build_oracle()
        self.lib = L = _load(ORACLE_SO)
        L.orc_resize.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.orc_border.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_blur.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_fast.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orc_fast_atan2.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_hamming.argtypes = [C.c_void_p, C.c_void_p]
        for f in (L.orc_knn2, L.orc_knn2_popcnt):
            f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_window_search.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 8
        L.orc_stereo_match.restype = C.c_int
        L.orc_stereo_match.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                       C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float,
                                       C.c_void_p, C.c_void_p, C.c_void_p]

    def search_by_projection(self, P, cur, last, taken=None, flags=None, octave=None, no_uright=False):
        """Oracle of ORBmatcher::SearchByProjection(Frame&, const Frame&, th, bMono); P is the same ProjParams struct
        the C ABI takes (layout-identical to the oracle's OrcProj)."""
        L = self.lib
        L.orc_search_by_projection.restype = C.c_int
        L.orc_search_by_projection.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [C.c_int] + [C.c_void_p] * 10
        if flags is None:
            flags = ((last.mp_valid != 0) & (last.mvbOutlier == 0)).astype(np.uint8) | ((last.mp_observed != 0).astype(np.uint8) << 1)
        flags = np.ascontiguousarray(flags, np.uint8)
        wp = np.ascontiguousarray(last.mp_world, np.float32)
        loct = np.ascontiguousarray(last.mvKeysUn["octave"] if octave is None else octave, np.int32)
        lang = np.ascontiguousarray(last.mvKeysUn["angle"], np.float32); ld = np.ascontiguousarray(last.mp_desc, np.uint8)
        kx = np.ascontiguousarray(cur.mvKeysUn["x"], np.float32); ky = np.ascontiguousarray(cur.mvKeysUn["y"], np.float32)
        koct = np.ascontiguousarray(cur.mvKeysUn["octave"], np.int32); kang = np.ascontiguousarray(cur.mvKeysUn["angle"], np.float32)
        ur = None if (cur.mvuRight is None or no_uright) else np.ascontiguousarray(cur.mvuRight, np.float32)
        tk = None if taken is None else np.ascontiguousarray(taken, np.uint8)
        cm = np.zeros(cur.N, np.int32); lm = np.zeros(last.N, np.int32); lk = np.zeros(last.N, np.uint8)
        n = L.orc_search_by_projection(C.addressof(P), last.N, _p(wp), _p(flags), _p(loct), _p(lang), _p(ld), cur.N, _p(kx),
                                       _p(ky), _p(koct), _p(kang), None if ur is None else _p(ur), _p(cur.mDescriptors),
                                       None if tk is None else _p(tk), _p(cm), _p(lm), _p(lk))
        return n, cm, lm, lk

    def search_local_points(self, P, nn_ratio, F, a, taken=None):
        """Oracle of ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>, th); `a` = ORBmatcher._mp_arrays(mps)."""
        L = self.lib
        L.orc_search_local_points.restype = C.c_int
        L.orc_search_local_points.argtypes = [C.c_void_p, C.c_float, C.c_int] + [C.c_void_p] * 7 + [C.c_int] + [C.c_void_p] * 8
        kx = np.ascontiguousarray(F.mvKeysUn["x"], np.float32); ky = np.ascontiguousarray(F.mvKeysUn["y"], np.float32)
        koct = np.ascontiguousarray(F.mvKeysUn["octave"], np.int32)
        ur = None if F.mvuRight is None else np.ascontiguousarray(F.mvuRight, np.float32)
        tk = None if taken is None else np.ascontiguousarray(taken, np.uint8)
        n_mp = len(a["flags"])
        cm = np.zeros(F.N, np.int32); mm = np.zeros(n_mp, np.int32)
        n = L.orc_search_local_points(C.addressof(P), float(nn_ratio), n_mp, _p(a["px"]), _p(a["py"]), _p(a["pxr"]),
                                      _p(a["lvl"]), _p(a["vc"]), _p(a["flags"]), _p(a["desc"]), F.N, _p(kx), _p(ky), _p(koct),
                                      None if ur is None else _p(ur), _p(F.mDescriptors), None if tk is None else _p(tk),
                                      _p(cm), _p(mm))
        return n, cm, mm

    def distinctive_descriptors(self, desc, off):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); off = np.ascontiguousarray(off, np.int32)
        n = len(off) - 1
        bi = np.zeros(n, np.int32); bm = np.zeros(n, np.int32)
        self.lib.orc_distinctive_descriptors.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        self.lib.orc_distinctive_descriptors(_p(desc) if desc.size else None, _p(off), n, _p(bi), _p(bm))
        return bi, bm

    def resize(self, src, dw, dh):
        src = np.ascontiguousarray(src, np.uint8); out = np.zeros((dh, dw), np.uint8)
        self.lib.orc_resize(_p(src), src.shape[1], src.shape[0], _p(out), dw, dh)
        return out

    def border(self, src, e):
        src = np.ascontiguousarray(src, np.uint8)
        out = np.zeros((src.shape[0] + 2 * e, src.shape[1] + 2 * e), np.uint8)
        self.lib.orc_border(_p(src), src.shape[1], src.shape[0], _p(out), e)
        return out

    def blur(self, src):
        src = np.ascontiguousarray(src, np.uint8); out = np.zeros_like(src)
        self.lib.orc_blur(_p(src), src.shape[1], src.shape[0], _p(out))
        return out

    def fast(self, img, th):
        img = np.ascontiguousarray(img, np.uint8)
        cap = img.size // 4 + 16
        out = np.zeros((cap, 3), np.int32)
        n = self.lib.orc_fast(_p(img), img.shape[1], img.shape[0], th, _p(out), cap)
        return out[:n].copy()

    def fast_atan2(self, y, x):
        y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
        out = np.zeros_like(y)
        self.lib.orc_fast_atan2(_p(y), _p(x), _p(out), y.size)
        return out

    def hamming(self, a, b):
        a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
        return self.lib.orc_hamming(_p(a), _p(b))

    def knn2(self, q, t, popcnt=False):
        q = np.ascontiguousarray(q, np.uint8); t = np.ascontiguousarray(t, np.uint8)
        i1 = np.zeros(len(q), np.int32); d1 = np.zeros(len(q), np.uint16); d2 = np.zeros(len(q), np.uint16)
        (self.lib.orc_knn2_popcnt if popcnt else self.lib.orc_knn2)(_p(q), len(q), _p(t), len(t), _p(i1), _p(d1), _p(d2))
        return i1, d1, d2

    def window_search(self, q, off, idx, skip, t):
        q = np.ascontiguousarray(q, np.uint8); t = np.ascontiguousarray(t, np.uint8)
        off = np.ascontiguousarray(off, np.int32); idx = np.ascontiguousarray(idx, np.int32)
        nq = len(q)
        bi = np.zeros(nq, np.int32); bd = np.zeros(nq, np.uint16); sd = np.zeros(nq, np.uint16); si = np.zeros(nq, np.int32)
        sk = None if skip is None else np.ascontiguousarray(skip, np.uint8)
        self.lib.orc_window_search(_p(q), nq, _p(off), _p(idx), None if sk is None else _p(sk), _p(t),
                                   _p(bi), _p(bd), _p(sd), _p(si))
        return bi, bd, sd, si

    def stereo_match(self, kL, dL, kR, dR, pyrL, pyrR, scale, inv_scale, mbf, minD, maxD):
        """pyrL/pyrR: lists of contiguous 2-D uint8 level interiors (un-blurred)."""
        nl = len(pyrL)
        PL = (C.c_void_p * nl)(*[p.ctypes.data for p in pyrL]); PR = (C.c_void_p * nl)(*[p.ctypes.data for p in pyrR])
        lw = np.array([p.shape[1] for p in pyrL], np.int32); lh = np.array([p.shape[0] for p in pyrL], np.int32)
        pl = np.array([p.strides[0] for p in pyrL], np.int32); pr = np.array([p.strides[0] for p in pyrR], np.int32)
        sc = np.ascontiguousarray(scale, np.float32); isc = np.ascontiguousarray(inv_scale, np.float32)
        NL = len(kL)
        ur = np.zeros(NL, np.float32); dp = np.zeros(NL, np.float32); sad = np.zeros(NL, np.int32)
        kL = np.ascontiguousarray(kL); kR = np.ascontiguousarray(kR)
        dL = np.ascontiguousarray(dL, np.uint8); dR = np.ascontiguousarray(dR, np.uint8)
        n = self.lib.orc_stereo_match(_p(kL), NL, _p(dL), _p(kR), len(kR), _p(dR), PL, PR, _p(lw), _p(lh), _p(pl), _p(pr),
                                      nl, _p(sc), _p(isc), mbf, minD, maxD, _p(ur), _p(dp), _p(sad))
        return ur, dp, sad, n


def canonical(kps, desc):
    """Sort (level, y, x) — the order-free comparison key of SURVEY.md §7 H8."""
    order = np.lexsort((kps["x"], kps["y"], kps["octave"]))
    return kps[order], desc[order]


REF_FBOW_SO = os.path.join(ROOT, "oracle", "_ref", "libref_fbow.so")


def have_ref_fbow():
    build_oracle()
    return os.path.exists(REF_FBOW_SO)


def _bow_buffers(n):
    return (np.zeros(n, np.uint32), np.zeros(n, np.float32), np.zeros(n, np.uint32), np.zeros(n + 1, np.int32),
            np.zeros(n, np.uint32))


def _bow_pack(wid, ww, nw, nid, noff, feat, nn):
    """(fBow as (ids, weights), fBow2 as {node: feature indices})"""
    return (wid[:nw].copy(), ww[:nw].copy()), {int(nid[i]): feat[noff[i]:noff[i + 1]].copy() for i in range(nn)}


class OracleVocabulary:
    """Restatement of fbow::Vocabulary::transform (oracle/orb_oracle.cpp orc_fbow_*) over the bytes of a .fbow file."""

    def __init__(self, blob):
        build_oracle()
        self.L = L = _load(ORACLE_SO)
        self.blob = bytes(blob)
        L.orc_fbow_transform.argtypes = [C.c_char_p, C.c_size_t, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 7
        L.orc_fbow_normalize.argtypes = [C.c_void_p, C.c_int]
        L.orc_fbow_score.restype = C.c_double
        L.orc_fbow_score.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]

    def transform(self, desc, level=None):
        """level given: Vocabulary::transform(features, level, fBow&, fBow2&) -> ((ids, weights), {node: features});
        level None: Vocabulary::transform(features) -> L2-normalised (ids, weights)."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        wid, ww, nid, noff, feat = _bow_buffers(n)
        nw, nn = C.c_int(), C.c_int()
        rc = self.L.orc_fbow_transform(self.blob, len(self.blob), _p(desc), n, -1 if level is None else level, _p(wid),
                                       _p(ww), C.byref(nw), _p(nid), _p(noff), _p(feat), C.byref(nn))
        assert rc == 0, "bad vocabulary blob"
        bow, fv = _bow_pack(wid, ww, nw.value, nid, noff, feat, nn.value)
        if level is None:
            self.L.orc_fbow_normalize(_p(bow[1]), len(bow[1]))
            return bow
        return bow, fv

    def score(self, a, b):
        return self.L.orc_fbow_score(_p(a[0]), _p(a[1]), len(a[0]), _p(b[0]), _p(b[1]), len(b[0]))


class RefVocabulary:
    """The unmodified reference fbow::Vocabulary (oracle/_ref/libref_fbow.so), loaded through fromStream."""

    def __init__(self, blob):
        build_oracle()
        self.L = L = _load(REF_FBOW_SO)
        L.ref_fbow_load.restype = C.c_void_p
        L.ref_fbow_load.argtypes = [C.c_char_p, C.c_size_t]
        L.ref_fbow_free.argtypes = [C.c_void_p]
        L.ref_fbow_info.argtypes = [C.c_void_p] * 5
        L.ref_fbow_transform2.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 7
        L.ref_fbow_transform.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 3
        L.ref_fbow_score.restype = C.c_double
        L.ref_fbow_score.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        self.h = L.ref_fbow_load(bytes(blob), len(blob))
        assert self.h, "reference fromStream rejected the blob"

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_fbow_free(self.h)
            self.h = None

    def info(self):
        k, nb, ds, tw = C.c_uint32(), C.c_uint64(), C.c_uint32(), C.c_uint64()
        self.L.ref_fbow_info(self.h, C.byref(k), C.byref(nb), C.byref(ds), C.byref(tw))
        return dict(k=k.value, nblocks=nb.value, desc_size=ds.value, total_words=tw.value)

    def transform(self, desc, level=None):
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        wid, ww, nid, noff, feat = _bow_buffers(n)
        nw, nn = C.c_int(), C.c_int()
        if level is None:
            rc = self.L.ref_fbow_transform(self.h, _p(desc), n, _p(wid), _p(ww), C.byref(nw))
            assert rc == 0
            return wid[:nw.value].copy(), ww[:nw.value].copy()
        rc = self.L.ref_fbow_transform2(self.h, _p(desc), n, level, _p(wid), _p(ww), C.byref(nw), _p(nid), _p(noff),
                                        _p(feat), C.byref(nn))
        assert rc == 0
        return _bow_pack(wid, ww, nw.value, nid, noff, feat, nn.value)

    def score(self, a, b):
        return self.L.ref_fbow_score(_p(a[0]), _p(a[1]), len(a[0]), _p(b[0]), _p(b[1]), len(b[0]))


class OraclePrep:
    """Upstream image prep restatements (oracle/cvprim.hpp): cvtColor X2GRAY, remap INTER_LINEAR, undistortPoints."""

    def __init__(self):
        build_oracle()
        self.L = L = _load(ORACLE_SO)
        L.orc_cvt_gray.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
        L.orc_remap.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.orc_undistort_points.argtypes = [C.c_void_p, C.c_int] + [C.c_float] * 4 + [C.c_void_p, C.c_int, C.c_void_p]

    def cvt_gray(self, img, blue_first):
        img = np.ascontiguousarray(img, np.uint8)
        h, w, ch = img.shape
        out = np.zeros((h, w), np.uint8)
        self.L.orc_cvt_gray(_p(img), w, h, w * ch, ch, int(blue_first), _p(out))
        return out

    def remap(self, src, mapx, mapy):
        src = np.ascontiguousarray(src, np.uint8)
        mapx, mapy = np.ascontiguousarray(mapx, np.float32), np.ascontiguousarray(mapy, np.float32)
        dh, dw = mapx.shape
        out = np.zeros((dh, dw), np.uint8)
        self.L.orc_remap(_p(src), src.shape[1], src.shape[0], _p(mapx), _p(mapy), dw, dh, _p(out))
        return out

    def undistort_points(self, pts, fx, fy, cx, cy, dist):
        pts = np.ascontiguousarray(pts, np.float32).reshape(-1, 2)
        dist = np.ascontiguousarray(dist, np.float32).ravel()
        out = np.zeros_like(pts)
        self.L.orc_undistort_points(_p(pts), len(pts), fx, fy, cx, cy, _p(dist), len(dist), _p(out))
        return out


def oracle_search_by_bow(a, b, nn_ratio, th, strict, check_ori):
    """orc_search_by_bow over two ORBmatcher.bow_side dicts."""
    build_oracle()
    L = _load(ORACLE_SO)
    L.orc_search_by_bow.restype = C.c_int
    side = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_search_by_bow.argtypes = side + side + [C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p]

    def args(d, need):
        e = d["eligible"]
        if e is None and need:
            e = np.ones(d["n"], np.uint8)
        return [d["n"], None if e is None else _p(e), _p(d["angle"]), _p(d["desc"]), len(d["node_id"]), _p(d["node_id"]),
                _p(d["node_off"]), _p(d["node_feat"])], e

    aa, ka = args(a, True)
    bb, kb = args(b, False)
    match = np.full(a["n"], -1, np.int32)
    n = L.orc_search_by_bow(*aa, *bb, np.float32(nn_ratio), th, int(strict), int(check_ori), _p(match))
    return n, match


def oracle_search_for_triangulation(a, b, geo, th, check_ori):
    """orc_search_for_triangulation over two bow_side dicts and the geometry dict of ORBmatcher.SearchForTriangulation."""
    build_oracle()
    L = _load(ORACLE_SO)
    L.orc_search_for_triangulation.restype = C.c_int
    side = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_search_for_triangulation.argtypes = side + side + [C.c_void_p] * 6 + [C.c_float, C.c_float, C.c_void_p, C.c_void_p,
                                                                                C.c_int, C.c_int, C.c_void_p]
    keep = []

    def args(d):
        e = d["eligible"] if d["eligible"] is not None else np.ones(d["n"], np.uint8)
        keep.append(e)
        return [d["n"], _p(e), _p(d["angle"]), _p(d["desc"]), len(d["node_id"]), _p(d["node_id"]), _p(d["node_off"]),
                _p(d["node_feat"])]

    g = {k: np.ascontiguousarray(geo[k], t) for k, t in (("pt_a", np.float32), ("pt_b", np.float32), ("octave_b", np.int32),
                                                          ("stereo_a", np.uint8), ("stereo_b", np.uint8), ("F12", np.float32),
                                                          ("scale_factors", np.float32), ("level_sigma2", np.float32))}
    match = np.full(a["n"], -1, np.int32)
    n = L.orc_search_for_triangulation(*args(a), *args(b), _p(g["pt_a"]), _p(g["pt_b"]), _p(g["octave_b"]), _p(g["stereo_a"]),
                                       _p(g["stereo_b"]), _p(g["F12"]), np.float32(geo["ex"]), np.float32(geo["ey"]),
                                       _p(g["scale_factors"]), _p(g["level_sigma2"]), th, int(check_ori), _p(match))
    return n, match


def oracle_fuse_search(P, inv_sigma2, chi2, pts, kf):
    """orc_fuse_search; P = the ProjParams the C ABI took (layout-identical to OrcProj)."""
    build_oracle()
    L = _load(ORACLE_SO)
    L.orc_fuse_search.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 6 + [C.c_int] + [C.c_void_p] * 7
    a = {k: np.ascontiguousarray(pts[k], t) for k, t in (("u", np.float32), ("v", np.float32), ("ur", np.float32),
                                                          ("level", np.int32), ("valid", np.uint8), ("desc", np.uint8))}
    n = len(a["u"])
    kx = np.ascontiguousarray(kf.mvKeysUn["x"], np.float32); ky = np.ascontiguousarray(kf.mvKeysUn["y"], np.float32)
    ko = np.ascontiguousarray(kf.mvKeysUn["octave"], np.int32)
    ur = None if kf.mvuRight is None else np.ascontiguousarray(kf.mvuRight, np.float32)
    bi, bd = np.full(n, -1, np.int32), np.full(n, 256, np.uint16)
    L.orc_fuse_search(C.addressof(P), _p(inv_sigma2), int(chi2), n, _p(a["u"]), _p(a["v"]), _p(a["ur"]), _p(a["level"]),
                      _p(a["valid"]), _p(a["desc"]), kf.N, _p(kx), _p(ky), _p(ko), None if ur is None else _p(ur),
                      _p(kf.mDescriptors), _p(bi), _p(bd))
    return bi, bd


def oracle_search_kf_projection(P, pts, kf, taken, th_low=50):
    """orc_search_kf_projection; P = the ProjParams the C ABI took (layout-identical to OrcProj)."""
    build_oracle()
    L = _load(ORACLE_SO)
    L.orc_search_kf_projection.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 5 + [C.c_int] + [C.c_void_p] * 5 + [C.c_int] + [C.c_void_p] * 2
    L.orc_search_kf_projection.restype = C.c_int
    a = {k: np.ascontiguousarray(pts[k], t) for k, t in (("u", np.float32), ("v", np.float32), ("level", np.int32),
                                                          ("valid", np.uint8), ("desc", np.uint8))}
    n = len(a["u"])
    kx = np.ascontiguousarray(kf.mvKeysUn["x"], np.float32); ky = np.ascontiguousarray(kf.mvKeysUn["y"], np.float32)
    ko = np.ascontiguousarray(kf.mvKeysUn["octave"], np.int32)
    tk = None if taken is None else np.ascontiguousarray(taken, np.uint8)
    km, pm = np.full(kf.N, -1, np.int32), np.full(n, -1, np.int32)
    r = L.orc_search_kf_projection(C.addressof(P), n, _p(a["u"]), _p(a["v"]), _p(a["level"]), _p(a["valid"]), _p(a["desc"]),
                                   kf.N, _p(kx), _p(ky), _p(ko), _p(kf.mDescriptors), None if tk is None else _p(tk),
                                   int(th_low), _p(km), _p(pm))
    return r, km, pm


def oracle_search_for_initialization(P, nn_ratio, F1, F2, prev, window_size, th_low=50):
    """orc_search_for_initialization; prev (F1.N, 2) float32 is updated in place like vbPrevMatched."""
    build_oracle()
    L = _load(ORACLE_SO)
    L.orc_search_for_initialization.argtypes = ([C.c_void_p, C.c_float, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 4 +
                                                [C.c_int] + [C.c_void_p] * 6)
    L.orc_search_for_initialization.restype = C.c_int
    o1 = np.ascontiguousarray(F1.mvKeysUn["octave"], np.int32); a1 = np.ascontiguousarray(F1.mvKeysUn["angle"], np.float32)
    kx = np.ascontiguousarray(F2.mvKeysUn["x"], np.float32); ky = np.ascontiguousarray(F2.mvKeysUn["y"], np.float32)
    ko = np.ascontiguousarray(F2.mvKeysUn["octave"], np.int32); ka = np.ascontiguousarray(F2.mvKeysUn["angle"], np.float32)
    assert prev.dtype == np.float32 and prev.flags.c_contiguous
    m12 = np.full(F1.N, -1, np.int32)
    r = L.orc_search_for_initialization(C.addressof(P), float(nn_ratio), int(th_low), int(window_size), F1.N, _p(o1), _p(a1),
                                        _p(F1.mDescriptors), _p(prev), F2.N, _p(kx), _p(ky), _p(ko), _p(ka),
                                        _p(F2.mDescriptors), _p(m12))
    return r, m12
