"""Array-of-fields mirror of the parts of ORB_SLAM2::Frame that the windowed searches read
(/root/reference/include/Frame.h:36-37, 110-190; src/Frame.cc:86-113 bounds and grid constants).

The reference keeps `std::vector<MapPoint*> mvpMapPoints`; here a frame's map points are parallel arrays
(world position, descriptor, Observations() > 0) because that is what crosses the C ABI."""
import numpy as np

from .extractor import KP_DTYPE

FRAME_GRID_COLS, FRAME_GRID_ROWS = 64, 48  # include/Frame.h:36-37


class Frame:
    def __init__(self, keys_un, descriptors, scale_factors, fx, fy, cx, cy, bf, width, height, u_right=None):
        self.mvKeysUn = np.ascontiguousarray(keys_un, KP_DTYPE)
        self.N = len(self.mvKeysUn)
        self.mDescriptors = np.ascontiguousarray(descriptors, np.uint8).reshape(self.N, 32)
        self.mvuRight = None if u_right is None else np.ascontiguousarray(u_right, np.float32)
        self.mvScaleFactors = np.ascontiguousarray(scale_factors, np.float32)
        self.fx, self.fy, self.cx, self.cy = (np.float32(v) for v in (fx, fy, cx, cy))
        self.mbf = np.float32(bf)
        self.mb = np.float32(self.mbf / self.fx)  # src/Frame.cc:106
        # no distortion: bounds are the image rectangle (src/Frame.cc:455-461)
        self.mnMinX, self.mnMaxX, self.mnMinY, self.mnMaxY = np.float32(0), np.float32(width), np.float32(0), np.float32(height)
        self.mfGridElementWidthInv = np.float32(np.float32(FRAME_GRID_COLS) / (self.mnMaxX - self.mnMinX))   # :101
        self.mfGridElementHeightInv = np.float32(np.float32(FRAME_GRID_ROWS) / (self.mnMaxY - self.mnMinY))  # :102
        self.mTcw = np.eye(4, dtype=np.float32)
        # map points (parallel arrays; index i <-> mvpMapPoints[i])
        self.mp_valid = np.zeros(self.N, np.uint8)        # mvpMapPoints[i] != NULL
        self.mvbOutlier = np.zeros(self.N, np.uint8)
        self.mp_observed = np.zeros(self.N, np.uint8)     # pMP->Observations() > 0
        self.mp_world = np.zeros((self.N, 3), np.float32)  # pMP->GetWorldPos()
        self.mp_desc = self.mDescriptors.copy()            # pMP->GetDescriptor()

    def SetPose(self, Tcw):
        self.mTcw = np.ascontiguousarray(Tcw, np.float32).reshape(4, 4)


def forward_backward(cur, last, bMono):
    """bForward / bBackward of src/ORBmatcher.cc:1334-1344 (float32 matrix algebra as cv::Mat CV_32F does it)."""
    Rcw, tcw = cur.mTcw[:3, :3], cur.mTcw[:3, 3]
    twc = -(Rcw.T.astype(np.float64) @ tcw.astype(np.float64)).astype(np.float32)
    Rlw, tlw = last.mTcw[:3, :3], last.mTcw[:3, 3]
    tlc = (Rlw.astype(np.float64) @ twc.astype(np.float64)).astype(np.float32) + tlw
    return bool(tlc[2] > cur.mb and not bMono), bool(-tlc[2] > cur.mb and not bMono)
