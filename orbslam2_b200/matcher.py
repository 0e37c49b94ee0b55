"""Host-side mirror of ORB_SLAM2::ORBmatcher's descriptor-distance surface (/root/reference/include/ORBmatcher.h:38-102)
over the C ABI: DescriptorDistance, the brute-force 2-NN + ratio test and the windowed best/second-best scan that every
ORBmatcher search loop runs. Geometric gating (projection, frustum, epipolar) stays with the caller,
who passes candidate lists. All compute happens in liborbx.so's CUDA kernels."""
import ctypes as C
import numpy as np

from . import _lib


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def _ptr(a):
    return a.data_ptr() if _is_torch(a) else a.ctypes.data


class ORBmatcher:
    TH_LOW = 50       # src/ORBmatcher.cc:35-37
    TH_HIGH = 100
    HISTO_LENGTH = 30

    def __init__(self, nnratio=0.6, checkOri=True, device=0):
        self.mfNNratio = np.float32(nnratio)
        self.mbCheckOrientation = checkOri
        self._L = _lib.lib()
        self._h = C.c_void_p()
        _lib.check(self._L.orbm_create(device, C.byref(self._h)))
        self.device = device
        self._space = (_lib.MEM_HOST, _lib.MEM_HOST)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbm_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _set_space(self, i, o):
        if (i, o) != self._space:
            _lib.check(self._L.orbm_set_memory_space(self._h, i, o))
            self._space = (i, o)

    def DescriptorDistance(self, a, b):
        """Hamming distance of two 32-byte descriptors (or of N row pairs -> array)."""
        a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32); b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
        d = np.zeros(len(a), np.uint16)
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(self._L.orbm_distance(self._h, a.ctypes.data, b.ctypes.data, len(a), d.ctypes.data))
        return int(d[0]) if len(d) == 1 else d.astype(np.int32)

    def set_train(self, t):
        dev = _is_torch(t)
        self._set_space(_lib.MEM_DEVICE if dev else _lib.MEM_HOST, self._space[1])
        if not dev:
            t = np.ascontiguousarray(t, np.uint8)
        self._train = t  # keep alive
        _lib.check(self._L.orbm_set_train(self._h, _ptr(t), int(t.shape[0])))

    def set_knn_mode(self, mode):
        """0 auto, 1 POPC kernel only, 2 tensor-core kernel (tcgen05) whenever a full 256-row train tile exists."""
        _lib.check(self._L.orbm_set_knn_mode(self._h, int(mode)))

    def last_knn_path(self):
        """1 = POPC kernel, 2 = tensor-core kernel (+ POPC remainder) for the last knn2 call."""
        return int(self._L.orbm_last_knn_path(self._h))

    def knn2(self, q, t=None):
        """(i1, d1, d2) per query; lowest train index wins ties; d2 counts multiplicity."""
        if _is_torch(q):
            import torch
            nq = q.shape[0]
            i1 = torch.empty(nq, dtype=torch.int32, device=q.device)
            d1 = torch.empty(nq, dtype=torch.int16, device=q.device); d2 = torch.empty(nq, dtype=torch.int16, device=q.device)
            self._set_space(_lib.MEM_DEVICE, _lib.MEM_DEVICE)
            torch.cuda.current_stream(q.device).synchronize()
            if t is not None:
                self._train = t
            _lib.check(self._L.orbm_knn2(self._h, q.data_ptr(), nq, 0 if t is None else t.data_ptr(),
                                         0 if t is None else t.shape[0], i1.data_ptr(), d1.data_ptr(), d2.data_ptr()))
            return i1, d1, d2
        q = np.ascontiguousarray(q, np.uint8)
        nq = len(q)
        i1 = np.zeros(nq, np.int32); d1 = np.zeros(nq, np.uint16); d2 = np.zeros(nq, np.uint16)
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        if t is not None:
            t = np.ascontiguousarray(t, np.uint8)
        _lib.check(self._L.orbm_knn2(self._h, q.ctypes.data, nq, None if t is None else t.ctypes.data,
                                     0 if t is None else len(t), i1.ctypes.data, d1.ctypes.data, d2.ctypes.data))
        return i1, d1, d2

    def match_ratio(self, q, t=None, th=None, ratio=None):
        """The SearchByFboW acceptance rule on a brute-force candidate set (src/ORBmatcher.cc:225-229):
        match iff d1 <= th and (float)d1 < ratio*(float)d2. Returns (i1 with -1 for rejects, d1, d2)."""
        th = self.TH_LOW if th is None else th
        ratio = self.mfNNratio if ratio is None else np.float32(ratio)
        i1, d1, d2 = self.knn2(q, t)
        ok = (d1 <= th) & (d1.astype(np.float32) < ratio * d2.astype(np.float32))
        return np.where(ok, i1, -1).astype(np.int32), d1, d2

    def window_search(self, q, cand_off, cand_idx, t=None, cand_skip=None):
        """Per query: scan candidates cand_idx[cand_off[a]:cand_off[a+1]] in list order ->
        (best_idx, best_d, second_d, second_idx); -1 / 256 when absent."""
        q = np.ascontiguousarray(q, np.uint8)
        off = np.ascontiguousarray(cand_off, np.int32); idx = np.ascontiguousarray(cand_idx, np.int32)
        sk = None if cand_skip is None else np.ascontiguousarray(cand_skip, np.uint8)
        nq = len(q)
        bi = np.zeros(nq, np.int32); si = np.zeros(nq, np.int32); bd = np.zeros(nq, np.uint16); sd = np.zeros(nq, np.uint16)
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        if t is not None:
            t = np.ascontiguousarray(t, np.uint8)
        _lib.check(self._L.orbm_window_search(self._h, q.ctypes.data, nq, off.ctypes.data, idx.ctypes.data if idx.size else None,
                                              None if sk is None else sk.ctypes.data, None if t is None else t.ctypes.data,
                                              0 if t is None else len(t), bi.ctypes.data, bd.ctypes.data, sd.ctypes.data,
                                              si.ctypes.data))
        return bi, bd, sd, si

    def _proj_params(self, cur, last, th, bMono):
        from .frame import forward_backward
        P = _lib.ProjParams()
        P.fx, P.fy, P.cx, P.cy, P.bf = float(cur.fx), float(cur.fy), float(cur.cx), float(cur.cy), float(cur.mbf)
        P.minX, P.maxX, P.minY, P.maxY = float(cur.mnMinX), float(cur.mnMaxX), float(cur.mnMinY), float(cur.mnMaxY)
        P.grid_w_inv, P.grid_h_inv = float(cur.mfGridElementWidthInv), float(cur.mfGridElementHeightInv)
        P.nlevels = len(cur.mvScaleFactors)
        for i, v in enumerate(cur.mvScaleFactors):
            P.scale[i] = float(v)
        for i, v in enumerate(cur.mTcw[:3, :3].reshape(-1)):
            P.Rcw[i] = float(v)
        for i, v in enumerate(cur.mTcw[:3, 3]):
            P.tcw[i] = float(v)
        fwd, bwd = forward_backward(cur, last, bMono)
        P.forward, P.backward = int(fwd), int(bwd)
        P.th, P.th_high, P.check_orientation = float(th), self.TH_HIGH, int(bool(self.mbCheckOrientation))
        return P

    def SearchByProjection(self, CurrentFrame, LastFrame, th, bMono, taken=None, device=False):
        """ORBmatcher::SearchByProjection(Frame &CurrentFrame, const Frame &LastFrame, th, bMono)
        (src/ORBmatcher.cc:1324-1466) entirely on the device: grid build, window candidates, Hamming scan, greedy claims
        in query order, rotation filter. Returns (nmatches, cur_match, last_match, last_keep); cur_match[i2] is the
        LastFrame index whose map point now sits in CurrentFrame.mvpMapPoints[i2] (-1 = none)."""
        cur, last = CurrentFrame, LastFrame
        P = self._proj_params(cur, last, th, bMono)
        flags = ((last.mp_valid != 0) & (last.mvbOutlier == 0)).astype(np.uint8) | ((last.mp_observed != 0).astype(np.uint8) << 1)
        keep = {"flags": flags, "oct": np.ascontiguousarray(last.mvKeysUn["octave"], np.int32),
                "ang": np.ascontiguousarray(last.mvKeysUn["angle"], np.float32),
                "wp": np.ascontiguousarray(last.mp_world, np.float32), "ld": np.ascontiguousarray(last.mp_desc, np.uint8),
                "tk": None if taken is None else np.ascontiguousarray(taken, np.uint8)}
        LA = _lib.LastFrameArgs(last.N, keep["wp"].ctypes.data, keep["flags"].ctypes.data, keep["oct"].ctypes.data,
                                keep["ang"].ctypes.data, keep["ld"].ctypes.data)
        CA = _lib.CurFrameArgs(cur.N, cur.mvKeysUn.ctypes.data, None if cur.mvuRight is None else cur.mvuRight.ctypes.data,
                               cur.mDescriptors.ctypes.data, None if keep["tk"] is None else keep["tk"].ctypes.data)
        n = C.c_int()
        if device:  # the same call with every array resident on the GPU (ORBX_MEM_DEVICE in and out)
            import torch
            up = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()
            ptr = lambda t: None if t is None else t.data_ptr()
            d = {k: up(v) for k, v in keep.items()}
            dk, dur, dd = up(cur.mvKeysUn), up(cur.mvuRight), up(cur.mDescriptors)
            LA = _lib.LastFrameArgs(last.N, ptr(d["wp"]), ptr(d["flags"]), ptr(d["oct"]), ptr(d["ang"]), ptr(d["ld"]))
            CA = _lib.CurFrameArgs(cur.N, ptr(dk), ptr(dur), ptr(dd), ptr(d["tk"]))
            tcm = torch.zeros(max(cur.N, 1), dtype=torch.int32, device="cuda")
            tlm = torch.zeros(max(last.N, 1), dtype=torch.int32, device="cuda")
            tlk = torch.zeros(max(last.N, 1), dtype=torch.uint8, device="cuda")
            torch.cuda.synchronize()
            self._set_space(_lib.MEM_DEVICE, _lib.MEM_DEVICE)
            _lib.check(self._L.orbm_search_by_projection(self._h, C.byref(P), C.byref(LA), C.byref(CA), tcm.data_ptr(),
                                                         tlm.data_ptr(), tlk.data_ptr(), C.byref(n)))
            self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
            return n.value, tcm[:cur.N].cpu().numpy(), tlm[:last.N].cpu().numpy(), tlk[:last.N].cpu().numpy()
        cm = np.zeros(cur.N, np.int32); lm = np.zeros(last.N, np.int32); lk = np.zeros(last.N, np.uint8)
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(self._L.orbm_search_by_projection(self._h, C.byref(P), C.byref(LA), C.byref(CA), cm.ctypes.data,
                                                     lm.ctypes.data, lk.ctypes.data, C.byref(n)))
        return n.value, cm, lm, lk

    def SearchByProjectionKF(self, CurrentFrame, kf, usable, predicted_level, th, ORBdist, taken=None):
        """ORBmatcher::SearchByProjection(Frame &CurrentFrame, KeyFrame *pKF, const set<MapPoint*> &sAlreadyFound, th,
        ORBdist) (src/ORBmatcher.cc:1468-1595, Tracking::Relocalization) on the device. `kf` is a Frame mirror of pKF
        (map-point positions, descriptors, mvKeysUn angles); usable[i] = map point present, not bad, not in sAlreadyFound
        and dist3D inside its scale-invariance range; predicted_level[i] = pMP->PredictScale(dist3D, &CurrentFrame)
        (both host side: a norm and a logarithm per point); taken[i2] = CurrentFrame.mvpMapPoints[i2] != NULL.
        Returns (nmatches, cur_match, kf_match, kf_keep) as SearchByProjection does."""
        cur = CurrentFrame
        P = self._proj_params(cur, kf, th, True)
        P.forward = P.backward = 0
        P.th_high, P.skip_depth_test = int(ORBdist), 1
        keep = {"flags": (np.ascontiguousarray(usable, np.uint8) != 0).astype(np.uint8) | np.uint8(2),
                "oct": np.ascontiguousarray(predicted_level, np.int32),
                "ang": np.ascontiguousarray(kf.mvKeysUn["angle"], np.float32),
                "wp": np.ascontiguousarray(kf.mp_world, np.float32), "ld": np.ascontiguousarray(kf.mp_desc, np.uint8),
                "tk": None if taken is None else np.ascontiguousarray(taken, np.uint8)}
        LA = _lib.LastFrameArgs(kf.N, keep["wp"].ctypes.data, keep["flags"].ctypes.data, keep["oct"].ctypes.data,
                                keep["ang"].ctypes.data, keep["ld"].ctypes.data)
        CA = _lib.CurFrameArgs(cur.N, cur.mvKeysUn.ctypes.data, None, cur.mDescriptors.ctypes.data,
                               None if keep["tk"] is None else keep["tk"].ctypes.data)
        cm = np.zeros(cur.N, np.int32); lm = np.zeros(kf.N, np.int32); lk = np.zeros(kf.N, np.uint8)
        n = C.c_int()
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(self._L.orbm_search_by_projection(self._h, C.byref(P), C.byref(LA), C.byref(CA), cm.ctypes.data,
                                                     lm.ctypes.data, lk.ctypes.data, C.byref(n)))
        return n.value, cm, lm, lk

    def FuseSearch(self, kf, pts, th=3.0, check_chi2=True):
        """Search half of ORBmatcher::Fuse (src/ORBmatcher.cc:821-971; check_chi2 False: the Sim3 overload :973-1096).
        kf: Frame mirror of the KeyFrame (mvKeysUn, mvuRight, mDescriptors, grid constants, mvScaleFactors); pts: dict
        of arrays u, v, ur, level, valid, desc for the map points the caller projected. Returns (best_idx, best_dist);
        the caller fuses where best_dist <= TH_LOW, in point order."""
        P = _lib.ProjParams()
        P.minX, P.maxX, P.minY, P.maxY = float(kf.mnMinX), float(kf.mnMaxX), float(kf.mnMinY), float(kf.mnMaxY)
        P.grid_w_inv, P.grid_h_inv = float(kf.mfGridElementWidthInv), float(kf.mfGridElementHeightInv)
        P.nlevels = len(kf.mvScaleFactors)
        for i, v in enumerate(kf.mvScaleFactors):
            P.scale[i] = float(v)
        P.th = float(th)
        inv_sigma2 = (np.float32(1.0) / (kf.mvScaleFactors * kf.mvScaleFactors)).astype(np.float32)
        a = {k: np.ascontiguousarray(pts[k], t) for k, t in (("u", np.float32), ("v", np.float32), ("ur", np.float32),
                                                              ("level", np.int32), ("valid", np.uint8), ("desc", np.uint8))}
        n = len(a["u"])
        FP = _lib.FusePoints(n, a["u"].ctypes.data, a["v"].ctypes.data, a["ur"].ctypes.data, a["level"].ctypes.data,
                             a["valid"].ctypes.data, a["desc"].ctypes.data)
        CA = _lib.CurFrameArgs(kf.N, kf.mvKeysUn.ctypes.data, None if kf.mvuRight is None else kf.mvuRight.ctypes.data,
                               kf.mDescriptors.ctypes.data, None)
        bi, bd = np.full(n, -1, np.int32), np.full(n, 256, np.uint16)
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(self._L.orbm_fuse_search(self._h, C.byref(P), inv_sigma2.ctypes.data, int(check_chi2), C.byref(FP),
                                            C.byref(CA), bi.ctypes.data, bd.ctypes.data))
        return bi, bd, P, inv_sigma2

    def _kf_grid_params(self, kf, th):
        P = _lib.ProjParams()
        P.minX, P.maxX, P.minY, P.maxY = float(kf.mnMinX), float(kf.mnMaxX), float(kf.mnMinY), float(kf.mnMaxY)
        P.grid_w_inv, P.grid_h_inv = float(kf.mfGridElementWidthInv), float(kf.mfGridElementHeightInv)
        P.nlevels = len(kf.mvScaleFactors)
        for i, v in enumerate(kf.mvScaleFactors):
            P.scale[i] = float(v)
        P.th = float(th)
        return P

    def SearchByProjectionSim3(self, kf, pts, th=10, taken=None):
        """ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, vpPoints, vpMatched, int th)
        (src/ORBmatcher.cc:285-398; LoopClosing::ComputeSim3) on the device. kf: Frame mirror of pKF; pts: dict of arrays
        u, v, level, valid, desc for the points the caller projected with Scw; taken[idx] = vpMatched[idx] != NULL on
        entry. Returns (nmatches, kf_match, pt_match, P): kf_match[idx] = point now in vpMatched[idx], -1 = unchanged."""
        P = self._kf_grid_params(kf, th)
        P.th_high = self.TH_LOW
        a = {k: np.ascontiguousarray(pts[k], t) for k, t in (("u", np.float32), ("v", np.float32), ("level", np.int32),
                                                              ("valid", np.uint8), ("desc", np.uint8))}
        n = len(a["u"])
        tk = None if taken is None else np.ascontiguousarray(taken, np.uint8)
        FP = _lib.FusePoints(n, a["u"].ctypes.data, a["v"].ctypes.data, None, a["level"].ctypes.data,
                             a["valid"].ctypes.data, a["desc"].ctypes.data)
        CA = _lib.CurFrameArgs(kf.N, kf.mvKeysUn.ctypes.data, None, kf.mDescriptors.ctypes.data,
                               None if tk is None else tk.ctypes.data)
        km, pm = np.full(kf.N, -1, np.int32), np.full(n, -1, np.int32)
        nm = C.c_int()
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(self._L.orbm_search_kf_projection(self._h, C.byref(P), C.byref(FP), C.byref(CA), km.ctypes.data,
                                                     pm.ctypes.data, C.byref(nm)))
        return nm.value, km, pm, P

    def SearchBySim3(self, kf1, kf2, pts1_in_2, pts2_in_1, th=7.5):
        """The two window searches of ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1098-1322) and the mutual-agreement
        step (:1303-1319). pts1_in_2: KF1's map points projected into KF2 with the Sim3 (dict u, v, level, valid, desc;
        entry i belongs to keypoint i of KF1, valid = has a point, not bad, not already matched, gates passed);
        pts2_in_1 likewise. Returns (nFound, match12) with match12[i1] = keypoint of KF2 agreed by both directions."""
        def one(kf, pts):
            p = dict(pts)
            p["ur"] = np.zeros(len(p["u"]), np.float32)
            bi, bd, _, _ = self.FuseSearch(kf, p, th=th, check_chi2=False)
            return np.where(bd <= self.TH_HIGH, bi, -1)
        m1 = one(kf2, pts1_in_2)
        m2 = one(kf1, pts2_in_1)
        i1 = np.arange(len(m1))
        ok = (m1 >= 0) & (m2[np.clip(m1, 0, max(len(m2) - 1, 0))] == i1) if len(m2) else np.zeros(len(m1), bool)
        match12 = np.where(ok, m1, -1).astype(np.int32)
        return int(ok.sum()), match12

    def SearchForInitialization(self, F1, F2, vbPrevMatched, windowSize=10):
        """ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, vbPrevMatched, vnMatches12, windowSize)
        (src/ORBmatcher.cc:400-515; Tracking::MonocularInitialization) on the device. vbPrevMatched: (F1.N, 2) float32,
        updated IN PLACE as the reference does. Returns (nmatches, vnMatches12, P)."""
        P = self._kf_grid_params(F2, float(windowSize))
        P.check_orientation = int(self.mbCheckOrientation)
        prev = vbPrevMatched
        assert prev.dtype == np.float32 and prev.shape == (F1.N, 2) and prev.flags.c_contiguous
        CA = _lib.CurFrameArgs(F2.N, F2.mvKeysUn.ctypes.data, None, F2.mDescriptors.ctypes.data, None)
        m12 = np.full(F1.N, -1, np.int32)
        nm = C.c_int()
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(self._L.orbm_search_for_initialization(self._h, C.byref(P), float(self.mfNNratio), int(self.TH_LOW),
                                                          int(windowSize), F1.N, F1.mvKeysUn.ctypes.data,
                                                          F1.mDescriptors.ctypes.data, prev.ctypes.data, C.byref(CA),
                                                          m12.ctypes.data, C.byref(nm)))
        return nm.value, m12, P

    def SearchLocalPoints(self, F, mps, th=1.0, taken=None):
        """ORBmatcher::SearchByProjection(Frame &F, const vector<MapPoint*> &vpMapPoints, th) (src/ORBmatcher.cc:43-127)
        on the device. `mps` is a dict of arrays: proj_x, proj_y, proj_xr, level, view_cos, in_view (mbTrackInView &&
        !isBad()), observed (Observations() > 0), desc [n,32]. Returns (nmatches, cur_match, mp_match)."""
        P = _lib.ProjParams()
        P.minX, P.maxX, P.minY, P.maxY = float(F.mnMinX), float(F.mnMaxX), float(F.mnMinY), float(F.mnMaxY)
        P.grid_w_inv, P.grid_h_inv = float(F.mfGridElementWidthInv), float(F.mfGridElementHeightInv)
        P.nlevels = len(F.mvScaleFactors)
        for i, v in enumerate(F.mvScaleFactors):
            P.scale[i] = float(v)
        P.th, P.th_high = float(th), self.TH_HIGH
        a = self._mp_arrays(mps)
        tk = None if taken is None else np.ascontiguousarray(taken, np.uint8)
        MA = _lib.MapPointsArgs(len(a["flags"]), a["px"].ctypes.data, a["py"].ctypes.data, a["pxr"].ctypes.data,
                                a["lvl"].ctypes.data, a["vc"].ctypes.data, a["flags"].ctypes.data, a["desc"].ctypes.data)
        CA = _lib.CurFrameArgs(F.N, F.mvKeysUn.ctypes.data, None if F.mvuRight is None else F.mvuRight.ctypes.data,
                               F.mDescriptors.ctypes.data, None if tk is None else tk.ctypes.data)
        cm = np.zeros(F.N, np.int32); mm = np.zeros(len(a["flags"]), np.int32)
        n = C.c_int()
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(self._L.orbm_search_local_points(self._h, C.byref(P), float(self.mfNNratio), C.byref(MA), C.byref(CA),
                                                    cm.ctypes.data, mm.ctypes.data, C.byref(n)))
        return n.value, cm, mm, P

    # --- SearchByFboW (src/ORBmatcher.cc:157-283 KeyFrame vs Frame, :517-650 KeyFrame vs KeyFrame) -----------------------
    @staticmethod
    def bow_side(desc, angle, feat_vec, eligible=None):
        """One frame for SearchByFboW: descriptors [n,32], keypoint angles [n], mFbowFeatVec as Vocabulary.transform returns
        it ({node: feature indices}), and the per-feature "has a good map point" mask (None: every feature)."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        keys = sorted(feat_vec)
        off = np.zeros(len(keys) + 1, np.int32)
        off[1:] = np.cumsum([len(feat_vec[k]) for k in keys])
        feat = (np.concatenate([np.asarray(feat_vec[k], np.uint32) for k in keys]) if keys else np.zeros(0, np.uint32))
        return dict(n=len(desc), desc=desc, angle=np.ascontiguousarray(angle, np.float32),
                    eligible=None if eligible is None else np.ascontiguousarray(eligible, np.uint8),
                    node_id=np.array(keys, np.uint32), node_off=off, node_feat=np.ascontiguousarray(feat, np.uint32))

    def SearchByFboW(self, a, b, strict=False, th=None, device=False):
        """a: KeyFrame (queries; eligible = map point present and not bad), b: Frame (strict False) or second KeyFrame
        (strict True, eligible = map point present and not bad). Returns (nmatches, match_a [a.n] int32: index in b or -1)."""
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        keep = []

        def side(d, need_elig):
            elig = d["eligible"]
            if elig is None and need_elig:
                elig = np.ones(d["n"], np.uint8)
            keep.append(elig)
            return _lib.BowSide(d["n"], None if elig is None else elig.ctypes.data, d["angle"].ctypes.data,
                                d["desc"].ctypes.data, len(d["node_id"]), d["node_id"].ctypes.data,
                                d["node_off"].ctypes.data, d["node_feat"].ctypes.data)

        n = C.c_int()
        if device:  # every array resident on the GPU (ORBX_MEM_DEVICE in and out)
            import torch
            up = lambda x: None if x is None else torch.from_numpy(np.ascontiguousarray(x).view(np.uint8).reshape(-1).copy()).cuda()
            ptr = lambda t: None if t is None else t.data_ptr()

            def dside(d, need_elig):
                elig = d["eligible"]
                if elig is None and need_elig:
                    elig = np.ones(d["n"], np.uint8)
                t = [up(elig), up(d["angle"]), up(d["desc"]), up(d["node_id"]), up(d["node_off"]), up(d["node_feat"])]
                keep.append(t)
                return _lib.BowSide(d["n"], ptr(t[0]), ptr(t[1]), ptr(t[2]), len(d["node_id"]), ptr(t[3]), ptr(t[4]), ptr(t[5]))

            A, B = dside(a, True), dside(b, False)
            tm = torch.full((max(a["n"], 1),), -1, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            self._set_space(_lib.MEM_DEVICE, _lib.MEM_DEVICE)
            _lib.check(self._L.orbm_search_by_bow(self._h, C.byref(A), C.byref(B), int(a["node_off"][-1]), int(b["node_off"][-1]),
                                                  float(self.mfNNratio), int(self.TH_LOW if th is None else th), int(strict),
                                                  int(self.mbCheckOrientation), tm.data_ptr(), C.byref(n)))
            self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
            return n.value, tm[:a["n"]].cpu().numpy()
        A, B = side(a, True), side(b, False)
        match = np.full(a["n"], -1, np.int32)
        _lib.check(self._L.orbm_search_by_bow(self._h, C.byref(A), C.byref(B), int(a["node_off"][-1]), int(b["node_off"][-1]),
                                              float(self.mfNNratio), int(self.TH_LOW if th is None else th), int(strict),
                                              int(self.mbCheckOrientation), match.ctypes.data, C.byref(n)))
        return n.value, match

    def SearchByFboWBatch(self, pairs, strict=False, th=None):
        """SearchByFboW for a list of (a, b) pairs in one launch sequence (orbm_search_by_bow_batch): the candidate
        keyframes of Relocalization / LoopClosing. pairs: [(a, b), ...] bow_side dicts as for SearchByFboW; the same b
        dict may appear in every pair (it is staged once). Returns [(nmatches, match_a), ...]."""
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        K = len(pairs)
        keep = []

        def side(d, need_elig):
            elig = d["eligible"]
            if elig is None and need_elig:
                elig = d.setdefault("_ones", np.ones(d["n"], np.uint8))
            keep.append(elig)
            return _lib.BowSide(d["n"], None if elig is None else elig.ctypes.data, d["angle"].ctypes.data,
                                d["desc"].ctypes.data, len(d["node_id"]), d["node_id"].ctypes.data,
                                d["node_off"].ctypes.data, d["node_feat"].ctypes.data)

        A = (_lib.BowSide * max(K, 1))(*[side(a, True) for a, _ in pairs])
        B = (_lib.BowSide * max(K, 1))(*[side(b, False) for _, b in pairs])
        npa = np.array([int(a["node_off"][-1]) for a, _ in pairs], np.int32)
        npb = np.array([int(b["node_off"][-1]) for _, b in pairs], np.int32)
        out = [np.full(a["n"], -1, np.int32) for a, _ in pairs]
        ptrs = (C.c_void_p * max(K, 1))(*[o.ctypes.data for o in out])
        nm = np.zeros(max(K, 1), np.int32)
        _lib.check(self._L.orbm_search_by_bow_batch(self._h, K, C.cast(A, C.c_void_p), C.cast(B, C.c_void_p), npa.ctypes.data,
                                                    npb.ctypes.data, float(self.mfNNratio),
                                                    int(self.TH_LOW if th is None else th), int(strict),
                                                    int(self.mbCheckOrientation), C.cast(ptrs, C.c_void_p), nm.ctypes.data))
        return [(int(nm[k]), out[k]) for k in range(K)]

    @staticmethod
    def _triangulation(geo, T=None):
        """orbm_triangulation from a geo dict; returns (struct, arrays to keep alive)."""
        g = {k: np.ascontiguousarray(geo[k], t) for k, t in (("pt_a", np.float32), ("pt_b", np.float32), ("octave_b", np.int32),
                                                              ("stereo_a", np.uint8), ("stereo_b", np.uint8))}
        T = _lib.Triangulation() if T is None else T
        T.pt_a, T.pt_b, T.octave_b = g["pt_a"].ctypes.data, g["pt_b"].ctypes.data, g["octave_b"].ctypes.data
        T.stereo_a, T.stereo_b = g["stereo_a"].ctypes.data, g["stereo_b"].ctypes.data
        F = np.ascontiguousarray(geo["F12"], np.float32).ravel()
        for i in range(9):
            T.F12[i] = F[i]
        T.ex, T.ey = float(np.float32(geo["ex"])), float(np.float32(geo["ey"]))
        sf, s2 = np.asarray(geo["scale_factors"], np.float32), np.asarray(geo["level_sigma2"], np.float32)
        T.nlevels = len(sf)
        for i in range(len(sf)):
            T.scale_factors[i], T.level_sigma2[i] = sf[i], s2[i]
        return T, g

    def SearchForTriangulationBatch(self, items, th=None):
        """SearchForTriangulation for a list of (a, b, geo) in one launch sequence (orbm_search_for_triangulation_batch):
        the new keyframe against its covisible neighbours in LocalMapping::CreateNewMapPoints. Arrays shared between the
        items (the new keyframe's side) are staged once. Returns [(nmatches, vMatches12), ...]."""
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        K = len(items)
        keep = []

        def side(d):
            e = d["eligible"] if d["eligible"] is not None else d.setdefault("_ones", np.ones(d["n"], np.uint8))
            keep.append(e)
            return _lib.BowSide(d["n"], e.ctypes.data, d["angle"].ctypes.data, d["desc"].ctypes.data, len(d["node_id"]),
                                d["node_id"].ctypes.data, d["node_off"].ctypes.data, d["node_feat"].ctypes.data)

        A = (_lib.BowSide * max(K, 1))(*[side(a) for a, _, _ in items])
        B = (_lib.BowSide * max(K, 1))(*[side(b) for _, b, _ in items])
        T = (_lib.Triangulation * max(K, 1))()
        for k, (_, _, geo) in enumerate(items):
            keep.append(self._triangulation(geo, T[k])[1])
        npa = np.array([int(a["node_off"][-1]) for a, _, _ in items], np.int32)
        npb = np.array([int(b["node_off"][-1]) for _, b, _ in items], np.int32)
        out = [np.full(a["n"], -1, np.int32) for a, _, _ in items]
        ptrs = (C.c_void_p * max(K, 1))(*[o.ctypes.data for o in out])
        nm = np.zeros(max(K, 1), np.int32)
        _lib.check(self._L.orbm_search_for_triangulation_batch(self._h, K, C.cast(A, C.c_void_p), C.cast(B, C.c_void_p),
                                                               npa.ctypes.data, npb.ctypes.data, C.cast(T, C.c_void_p),
                                                               int(self.TH_LOW if th is None else th),
                                                               int(self.mbCheckOrientation), C.cast(ptrs, C.c_void_p),
                                                               nm.ctypes.data))
        return [(int(nm[k]), out[k]) for k in range(K)]

    def SearchForTriangulation(self, a, b, geo, th=None):
        """SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:652-819). a, b: bow_side
        dicts with eligible = "no map point (and stereo when bOnlyStereo)"; geo: dict(pt_a [nA,2], pt_b [nB,2], octave_b,
        stereo_a, stereo_b, F12 [3,3], ex, ey, scale_factors, level_sigma2). Returns (nmatches, vMatches12)."""
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        ea = a["eligible"] if a["eligible"] is not None else np.ones(a["n"], np.uint8)
        eb = b["eligible"] if b["eligible"] is not None else np.ones(b["n"], np.uint8)

        def side(d, e):
            return _lib.BowSide(d["n"], e.ctypes.data, d["angle"].ctypes.data, d["desc"].ctypes.data, len(d["node_id"]),
                                d["node_id"].ctypes.data, d["node_off"].ctypes.data, d["node_feat"].ctypes.data)

        A, B = side(a, ea), side(b, eb)
        T, g = self._triangulation(geo)
        match = np.full(a["n"], -1, np.int32)
        n = C.c_int()
        _lib.check(self._L.orbm_search_for_triangulation(self._h, C.byref(A), C.byref(B), int(a["node_off"][-1]),
                                                         int(b["node_off"][-1]), C.byref(T),
                                                         int(self.TH_LOW if th is None else th),
                                                         int(self.mbCheckOrientation), match.ctypes.data, C.byref(n)))
        return n.value, match

    @staticmethod
    def _mp_arrays(mps):
        f32 = lambda k: np.ascontiguousarray(mps[k], np.float32)
        flags = (np.asarray(mps["in_view"]) != 0).astype(np.uint8) | ((np.asarray(mps["observed"]) != 0).astype(np.uint8) << 1)
        return {"px": f32("proj_x"), "py": f32("proj_y"), "pxr": f32("proj_xr"), "vc": f32("view_cos"),
                "lvl": np.ascontiguousarray(mps["level"], np.int32), "flags": flags,
                "desc": np.ascontiguousarray(mps["desc"], np.uint8)}

    def ComputeDistinctiveDescriptors(self, desc, off):
        """Batched MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-307): map point p owns descriptor rows
        off[p]:off[p+1]. Returns (best_idx, best_median) per point (best_idx = -1 without observations)."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); off = np.ascontiguousarray(off, np.int32)
        n = len(off) - 1
        bi = np.zeros(n, np.int32); bm = np.zeros(n, np.int32)
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(self._L.orbm_distinctive_descriptors(self._h, desc.ctypes.data if desc.size else None, off.ctypes.data, n,
                                                        bi.ctypes.data, bm.ctypes.data))
        return bi, bm

    def search_stats(self):
        r, c = C.c_int(), C.c_int64()
        _lib.check(self._L.orbm_search_stats(self._h, C.byref(r), C.byref(c)))
        return r.value, c.value

    def cuda_stream(self):
        s = C.c_void_p()
        _lib.check(self._L.orbm_stream(self._h, C.byref(s)))
        return s.value

    def launch_count(self):
        return int(self._L.orbm_launch_count(self._h))
