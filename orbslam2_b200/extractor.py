"""Host-side mirror of ORB_SLAM2::ORBextractor (/root/reference/include/ORBextractor.h:45-112) over the C ABI.

Same constructor arguments, same call shape (`extractor(image, mask) -> keypoints, descriptors`), same getters and the
`mvImagePyramid` member; plus a batched call and device-memory I/O for torch CUDA tensors. All compute happens in
liborbx.so's CUDA kernels — this file only marshals pointers.
"""
import ctypes as C
import numpy as np

from . import _lib

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])


def _is_torch(x):
    return type(x).__module__.startswith("torch")


class ORBextractor:
    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, patchSize=31, halfPatchSize=15,
                 edgeThreshold=19, device=0):
        # the reference's yaml defaults: 0 means "use 31/15/19" (src/Tracking.cc:118-123)
        patchSize = patchSize or 31
        halfPatchSize = halfPatchSize or 15
        edgeThreshold = edgeThreshold or 19
        self._L = _lib.lib()
        self.params = _lib.OrbxParams(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, patchSize, halfPatchSize,
                                      edgeThreshold)
        self._h = C.c_void_p()
        _lib.check(self._L.orbx_create(C.byref(self.params), device, C.byref(self._h)))
        self.device = device
        n = nlevels
        t = [np.zeros(n, np.float32) for _ in range(4)]
        self.mnFeaturesPerLevel = np.zeros(n, np.int32)
        self.umax = np.zeros(halfPatchSize + 1, np.int32)
        _lib.check(self._L.orbx_tables(self._h, *[a.ctypes.data for a in t], self.mnFeaturesPerLevel.ctypes.data,
                                       self.umax.ctypes.data))
        self.mvScaleFactor, self.mvInvScaleFactor, self.mvLevelSigma2, self.mvInvLevelSigma2 = t
        self._in_space = self._out_space = _lib.MEM_HOST
        self._last = None
        self._pending = None
        self._channels = 1
        self._out_size = None

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._L.orbx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- getters of the reference class ---------------------------------------------------------------------------
    def GetLevels(self):
        return self.params.nlevels

    def GetScaleFactor(self):
        return float(np.float64(np.float32(self.params.scaleFactor)))

    def GetScaleFactors(self):
        return self.mvScaleFactor.copy()

    def GetInverseScaleFactors(self):
        return self.mvInvScaleFactor.copy()

    def GetScaleSigmaSquares(self):
        return self.mvLevelSigma2.copy()

    def GetInverseScaleSigmaSquares(self):
        return self.mvInvLevelSigma2.copy()

    # --- extraction ------------------------------------------------------------------------------------------------
    def max_keypoints(self, w, h):
        cap = C.c_int()
        _lib.check(self._L.orbx_max_keypoints(self._h, w, h, C.byref(cap)))
        return cap.value

    def set_chunk(self, frames):
        """Frames per pipeline chunk of the host-buffer path (0 = library default)."""
        _lib.check(self._L.orbx_set_chunk(self._h, int(frames)))

    # --- upstream image prep fused into the ingest (cvtColor, remap; src/Tracking.cc:269-294, stereo_euroc.cc:136) -----
    _FORMATS = {"gray": _lib.FMT_GRAY8, "rgb": _lib.FMT_RGB8, "bgr": _lib.FMT_BGR8, "rgba": _lib.FMT_RGBA8,
                "bgra": _lib.FMT_BGRA8}

    def set_input_format(self, fmt):
        """'gray' (default), 'rgb', 'bgr', 'rgba', 'bgra': images are then [.., H, W, C] and are converted like
        cv::cvtColor(.., COLOR_x2GRAY) on the device."""
        _lib.check(self._L.orbx_set_input_format(self._h, self._FORMATS[fmt]))
        self._channels = 1 if fmt == "gray" else (3 if fmt in ("rgb", "bgr") else 4)

    def set_rectify_maps(self, map_x, map_y):
        """Maps of cv::initUndistortRectifyMap(.., CV_32F): every image is remapped (INTER_LINEAR) on the device before
        extraction; None, None switches it off."""
        if map_x is None:
            _lib.check(self._L.orbx_set_rectify_maps(self._h, None, None, 0, 0))
            self._out_size = None
            return
        mx, my = np.ascontiguousarray(map_x, np.float32), np.ascontiguousarray(map_y, np.float32)
        assert mx.ndim == 2 and mx.shape == my.shape
        _lib.check(self._L.orbx_set_rectify_maps(self._h, mx.ctypes.data, my.ctypes.data, mx.shape[1], mx.shape[0]))
        self._out_size = (mx.shape[1], mx.shape[0])

    def undistort_keypoints(self, K4, dist, cap=None):
        """Frame::UndistortKeyPoints for the last batch: returns mvKeysUn as [B, cap] KP_DTYPE (host)."""
        B, W, H = self._last
        cap = cap or self.max_keypoints(W, H)
        K4 = np.ascontiguousarray(K4, np.float32)
        dist = np.ascontiguousarray(dist, np.float32).ravel()
        out = np.zeros((B, cap), KP_DTYPE)
        self._set_space(self._in_space, _lib.MEM_HOST)
        _lib.check(self._L.orbx_undistort_keypoints(self._h, K4.ctypes.data, dist.ctypes.data if len(dist) else None,
                                                    len(dist), out.ctypes.data, cap))
        return out

    def _set_space(self, i, o):
        if (i, o) != (self._in_space, self._out_space):
            _lib.check(self._L.orbx_set_memory_space(self._h, i, o))
            self._in_space, self._out_space = i, o

    def __call__(self, image, mask=None):
        """operator()(image, mask, keypoints, descriptors): mask is ignored, as in the reference
        (include/ORBextractor.h:57). Returns (keypoints [KP_DTYPE], descriptors [N,32] uint8); empty image -> empty."""
        image = np.asarray(image)
        if image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == (2 if self._channels == 1 else 3), "CV_8UC1 image required"
        k, d, n = self.extract_batch(image[None])
        return k[0, :n[0]].copy(), d[0, :n[0]].copy()

    def extract_batch(self, images, out=None):
        """images: [B,H,W] uint8 ([B,H,W,C] with a colour input format) numpy array (host, any row stride) or torch CUDA
        tensor (device). Returns (kps [B,cap], desc [B,cap,32], n [B]) in the same memory space as the input (n always
        numpy)."""
        if _is_torch(images):
            return self._extract_batch_torch(images, out)
        C_ = self._channels
        assert images.dtype == np.uint8 and images.ndim == (3 if C_ == 1 else 4)
        if C_ > 1:
            assert images.shape[3] == C_
            images = np.ascontiguousarray(images)
            B, H, W = images.shape[:3]
            row, frame = W * C_, H * W * C_
        else:
            B, H, W = images.shape
            if images.strides[2] != 1 or images.strides[1] < W or (B > 1 and images.strides[0] < images.strides[1] * H):
                images = np.ascontiguousarray(images)
            row, frame = images.strides[1], images.strides[0]
        oW, oH = self._out_size or (W, H)
        cap = self.max_keypoints(oW, oH)
        if out is None:
            out = (np.zeros((B, cap), KP_DTYPE), np.zeros((B, cap, 32), np.uint8))
        kps, desc = out
        n = np.zeros(B, np.int32)
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(self._L.orbx_extract_batch(self._h, images.ctypes.data, B, W, H, row, frame,
                                              kps.ctypes.data, desc.ctypes.data, cap, n.ctypes.data))
        self._last = (B, oW, oH)
        return kps, desc, n

    def submit_batch(self, images, out):
        """First half of extract_batch (orbx_extract_batch_submit): enqueue and return. `images` [B,H,W] uint8 and
        `out` = (kps [B,cap] KP_DTYPE or [B,cap,7] float32, desc [B,cap,32] uint8) are numpy arrays or CPU torch tensors
        (pinned memory makes the copies asynchronous), or all torch CUDA tensors; they must stay alive until wait_batch().
        Two extractors submitted back to back run together on the device (left and right image of a stereo frame)."""
        C_ = self._channels
        if _is_torch(images) and images.is_cuda:  # device-resident batch: the same two halves with device pointers
            import torch
            kps, desc = out
            assert images.dtype == torch.uint8 and images.is_contiguous() and kps.is_cuda and desc.is_cuda
            B, H, W = images.shape[:3]
            cap = desc.shape[1]
            self._set_space(_lib.MEM_DEVICE, _lib.MEM_DEVICE)
            torch.cuda.current_stream(images.device).synchronize()
            _lib.check(self._L.orbx_extract_batch_submit(self._h, images.data_ptr(), B, W, H, W * C_, W * H * C_,
                                                         kps.data_ptr(), desc.data_ptr(), cap))
            self._pending = (images, out, np.zeros(B, np.int32))
            self._last = (B,) + (self._out_size or (W, H))
            return

        def view(a):
            return a.numpy() if _is_torch(a) else a
        images, kps, desc = view(images), view(out[0]), view(out[1])
        assert images.dtype == np.uint8 and images.ndim == (3 if C_ == 1 else 4) and images.strides[-1] == 1
        assert C_ == 1 or (images.shape[3] == C_ and images.strides[2] == C_)
        B, H, W = images.shape[:3]
        cap = desc.shape[1]
        assert desc.shape == (B, cap, 32) and kps.shape[:2] == (B, cap) and kps.nbytes == B * cap * 28
        self._set_space(_lib.MEM_HOST, _lib.MEM_HOST)
        _lib.check(self._L.orbx_extract_batch_submit(self._h, images.ctypes.data, B, W, H, images.strides[1],
                                                     images.strides[0], kps.ctypes.data, desc.ctypes.data, cap))
        self._pending = (images, out, np.zeros(B, np.int32))
        self._last = (B,) + (self._out_size or (W, H))

    def wait_batch(self):
        """Second half (orbx_extract_batch_wait): returns (kps, desc, n) of the submitted batch."""
        images, out, n = self._pending
        self._pending = None
        _lib.check(self._L.orbx_extract_batch_wait(self._h, n.ctypes.data))
        return out[0], out[1], n

    def _extract_batch_torch(self, images, out=None):
        import torch
        C_ = self._channels
        assert images.is_cuda and images.dtype == torch.uint8 and images.dim() == (3 if C_ == 1 else 4) and images.is_contiguous()
        assert images.device.index == self.device
        B, H, W = images.shape[:3]
        oW, oH = self._out_size or (W, H)
        cap = self.max_keypoints(oW, oH)
        if out is None:
            out = (torch.empty((B, cap, 7), dtype=torch.float32, device=images.device),
                   torch.empty((B, cap, 32), dtype=torch.uint8, device=images.device))
        kps, desc = out
        n = np.zeros(B, np.int32)
        self._set_space(_lib.MEM_DEVICE, _lib.MEM_DEVICE)
        torch.cuda.current_stream(images.device).synchronize()
        _lib.check(self._L.orbx_extract_batch(self._h, images.data_ptr(), B, W, H, W * C_, W * H * C_, kps.data_ptr(),
                                              desc.data_ptr(), cap, n.ctypes.data))
        self._last = (B, oW, oH)
        return kps, desc, n

    # --- mvImagePyramid ---------------------------------------------------------------------------------------------
    def level_size(self, level):
        w, h = C.c_int(), C.c_int()
        _lib.check(self._L.orbx_level_size(self._h, level, C.byref(w), C.byref(h)))
        return w.value, h.value

    def pyramid_level(self, level, frame=0, with_border=False):
        w, h = self.level_size(level)
        e = self.params.edgeThreshold if with_border else 0
        out = np.zeros((h + 2 * e, w + 2 * e), np.uint8)
        _lib.check(self._L.orbx_pyramid_level(self._h, frame, level, out.ctypes.data, out.strides[0], int(with_border)))
        return out

    @property
    def mvImagePyramid(self):
        """The reference's public member: one un-blurred image per level (interior ROI of a padded buffer whose
        19-px frame is BORDER_REFLECT_101, src/ORBextractor.cc:927-942). Fetched from the device on access."""
        e = self.params.edgeThreshold
        out = []
        for l in range(self.params.nlevels):
            padded = self.pyramid_level(l, 0, True)
            out.append(padded[e:-e, e:-e])
        return out

    def GetImagePyramid(self):
        return self.mvImagePyramid

    def blurred_level(self, level, frame=0):
        w, h = self.level_size(level)
        out = np.zeros((h, w), np.uint8)
        _lib.check(self._L.orbx_blurred_level(self._h, frame, level, out.ctypes.data, out.strides[0]))
        return out

    def candidates(self, level, frame=0):
        w, h = self.level_size(level)
        cap = w * h // 3 + 64
        out = np.zeros((cap, 3), np.int32)
        n = C.c_int()
        _lib.check(self._L.orbx_candidates(self._h, frame, level, out.ctypes.data, cap, C.byref(n)))
        return out[:n.value].copy()

    # --- instrumentation --------------------------------------------------------------------------------------------
    def cuda_stream(self):
        s = C.c_void_p()
        _lib.check(self._L.orbx_stream(self._h, C.byref(s)))
        return s.value

    def launch_count(self):
        return int(self._L.orbx_launch_count(self._h))

    def set_profiling(self, on):
        _lib.check(self._L.orbx_set_profiling(self._h, int(on)))

    def stage_ms(self):
        ms = np.zeros(7, np.float32)
        _lib.check(self._L.orbx_stage_ms(self._h, ms.ctypes.data))
        return dict(zip(["ingest", "pyramid", "fast", "quadtree", "blur", "orient_desc", "egress"], ms.tolist()))


def compute_stereo_matches(left, right, mbf, minD, maxD, frame=None):
    """Frame::ComputeStereoMatches (src/Frame.cc:464-642) on the state resident in two extractors that have just
    processed the left/right images. Returns (mvuRight, mvDepth) per left keypoint; frame=None -> all frames [B,cap]."""
    L = left._L
    B, W, H = left._last
    cap = left.max_keypoints(W, H)
    left._set_space(left._in_space, _lib.MEM_HOST)
    if frame is None:
        u = np.zeros((B, cap), np.float32); z = np.zeros((B, cap), np.float32)
        _lib.check(L.orbx_stereo_match_batch(left._h, right._h, mbf, minD, maxD, u.ctypes.data, z.ctypes.data, cap))
        return u, z
    u = np.zeros(cap, np.float32); z = np.zeros(cap, np.float32)
    _lib.check(L.orbx_stereo_match(left._h, right._h, frame, mbf, minD, maxD, u.ctypes.data, z.ctypes.data))
    return u, z


class ExtractorRing:
    """`depth` extractor handles used round-robin so that `depth` batches are in flight: while one batch's kernels run,
    the next one's images cross PCIe and the previous one's keypoints come back. push(images) submits a batch and
    returns the results of the batch pushed `depth` calls earlier (None while the ring fills); drain() returns the
    rest in order. Result arrays are the ring's own pinned buffers: they are overwritten `depth` pushes after the call
    that returned them."""

    def __init__(self, depth=2, chunk=96, device=0, **params):
        self.ex = [ORBextractor(**params, device=device) for _ in range(depth)]
        for e in self.ex:
            e.set_chunk(chunk)
        self._out = [None] * (2 * depth)  # two result sets per handle: a returned set stays valid for `depth` pushes
        self._flip = [0] * depth
        self._shape = None
        self._busy = [False] * depth
        self._at = 0

    def _buffers(self, k, B, cap):
        import torch
        if self._shape != (B, cap):  # (re)allocate every result set at once: pinning memory stalls the device
            for k2 in range(len(self.ex)):
                if self._busy[k2]:
                    raise RuntimeError("ExtractorRing: drain() before pushing a batch of a different shape")
            self._out = [(torch.empty((B, cap, 7), dtype=torch.float32).pin_memory(),
                          torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()) for _ in self._out]
            self._shape = (B, cap)
        self._flip[k] ^= 1
        return self._out[2 * k + self._flip[k]]

    def _finish(self, k):
        kps, desc, n = self.ex[k].wait_batch()
        self._busy[k] = False
        return kps.numpy().view(KP_DTYPE)[..., 0], desc.numpy(), n

    def push(self, images):
        k = self._at
        self._at = (k + 1) % len(self.ex)
        done = self._finish(k) if self._busy[k] else None
        B, H, W = images.shape[:3]
        oW, oH = self.ex[k]._out_size or (W, H)
        self.ex[k].submit_batch(images, self._buffers(k, B, self.ex[k].max_keypoints(oW, oH)))
        self._busy[k] = True
        return done

    def drain(self):
        res = []
        for i in range(len(self.ex)):
            k = (self._at + i) % len(self.ex)
            if self._busy[k]:
                res.append(self._finish(k))
        return res

    def launch_count(self):
        return sum(e.launch_count() for e in self.ex)

    def close(self):
        for e in self.ex:
            e.close()
