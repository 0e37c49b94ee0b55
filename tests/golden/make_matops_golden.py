"""Generates tests/golden/cv2_matops.npz — known answers from the real OpenCV (python cv2 4.13.0) for the CV_32F matrix
algebra that /root/reference/src/ORBmatcher.cc, KeyFrame.cc and Frame.cc evaluate through cv::Mat expressions:
  Rcw*x3Dw + tcw, sR21*p3Dc1 + t21      cv::gemm, flags 0, 3x3 * 3x1 (+ C)      src/ORBmatcher.cc:1356, 1153-1154, 313 ...
  -Rcw.t()*tcw                          cv::gemm, GEMM_1_T, alpha -1            src/ORBmatcher.cc:1334, 298, 1483
  -sR21*t12                             cv::gemm, flags 0, alpha -1             src/ORBmatcher.cc:1117
  cv::norm(PO), cv::norm(IL, IR, L1)    src/ORBmatcher.cc:330, src/Frame.cc:594
oracle/shim_matcher/cvmat_shim.hpp (the stand-in the unmodified reference is compiled against) must reproduce them
bit for bit (tests/test_cpu_refmatcher.py). Mat::dot and the Mat*scalar forms have no Python binding; the shim follows
OpenCV's sources for those (dotProd_32f: double accumulator; convertTo: src*(float)alpha).
Run in the build container:  python tests/golden/make_matops_golden.py
"""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
rng = np.random.default_rng(2024)
f32 = np.float32
out = {"cv2_version": np.array(cv2.__version__)}
T = 120


def mats(r, c, n=T, scale=10.0):
    return (rng.standard_normal((n, r, c)) * scale).astype(f32)


# op 0 / 1: plain products of the shapes the reference forms (3x3*3x1, 3x3*3x3, 4x4*4x4, 4x4*4x1, 1x3*3x3, 3x4*4x1)
for name, (m, k, n) in {"33_31": (3, 3, 1), "33_33": (3, 3, 3), "44_44": (4, 4, 4), "44_41": (4, 4, 1), "13_33": (1, 3, 3),
                        "34_41": (3, 4, 1)}.items():
    A, B, C = mats(m, k), mats(k, n), mats(m, n)
    out[f"mul_{name}_A"], out[f"mul_{name}_B"], out[f"mul_{name}_C"] = A, B, C
    out[f"mul_{name}_D"] = np.stack([cv2.gemm(A[i], B[i], 1.0, None, 0.0) for i in range(T)])
    out[f"mul_{name}_DC"] = np.stack([cv2.gemm(A[i], B[i], 1.0, C[i], 1.0) for i in range(T)])
    out[f"mul_{name}_DN"] = np.stack([cv2.gemm(A[i], B[i], -1.0, None, 0.0) for i in range(T)])
# op 2 / 3: transposed left operand
for name, (m, k, n) in {"33_31": (3, 3, 1), "33_33": (3, 3, 3)}.items():
    A, B = mats(k, m), mats(k, n)
    out[f"tmul_{name}_A"], out[f"tmul_{name}_B"] = A, B
    out[f"tmul_{name}_D"] = np.stack([cv2.gemm(A[i], B[i], 1.0, None, 0.0, flags=cv2.GEMM_1_T) for i in range(T)])
    out[f"tmul_{name}_DN"] = np.stack([cv2.gemm(A[i], B[i], -1.0, None, 0.0, flags=cv2.GEMM_1_T) for i in range(T)])
# norms
V = mats(3, 1)
out["norm_A"] = V
out["norm_L2"] = np.array([cv2.norm(V[i]) for i in range(T)], np.float64)
P, Q = (rng.integers(-255, 256, (T, 11, 11))).astype(f32), (rng.integers(-255, 256, (T, 11, 11))).astype(f32)
out["l1_A"], out["l1_B"] = P, Q
out["l1"] = np.array([cv2.norm(P[i], Q[i], cv2.NORM_L1) for i in range(T)], np.float64)
np.savez_compressed(os.path.join(HERE, "cv2_matops.npz"), **out)
print("wrote cv2_matops.npz", {k: v.shape for k, v in out.items() if k.endswith("_D")})
