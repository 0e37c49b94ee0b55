"""Generates tests/golden/cv2_kat.npz — known-answer vectors from the real OpenCV (python cv2 4.13.0, the only OpenCV
available offline) for every external primitive the reference's hot path calls (SURVEY.md §8 c):
  cv::resize INTER_LINEAR 8U        /root/reference/src/ORBextractor.cc:934
  cv::copyMakeBorder REFLECT_101    /root/reference/src/ORBextractor.cc:936-942
  cv::GaussianBlur 7x7 sigma 2      /root/reference/src/ORBextractor.cc:900
  cv::FAST(th, nonmax=true)         /root/reference/src/ORBextractor.cc:803,808
  cv::fastAtan2                     /root/reference/src/ORBextractor.cc:98
Run in the build container:  python tests/golden/make_golden.py   (cv2 does not exist on the GPU box; the .npz travels)
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from orbslam2_b200.synth import synth_image  # noqa: E402

cv2.setNumThreads(1)
out = {"cv2_version": np.array(cv2.__version__)}
rng = np.random.default_rng(123)
imgs = {
    "synth": synth_image(208, 160, 11),
    "noise": rng.integers(0, 256, (75, 97), dtype=np.uint8),
    "steps": (np.indices((64, 80)).sum(0) // 7 * 37 % 256).astype(np.uint8),
}
for name, img in imgs.items():
    out[f"img_{name}"] = img
    h, w = img.shape
    # pyramid-style chain: each level resized from the previous one, sizes as the extractor rounds them
    prev, sf = img, np.float32(1.0)
    for l in range(1, 4):
        sf = np.float32(np.float64(sf) * np.float64(np.float32(1.2)))
        dw, dh = int(np.rint(np.float32(w) / sf)), int(np.rint(np.float32(h) / sf))
        prev = cv2.resize(prev, (dw, dh), interpolation=cv2.INTER_LINEAR)
        out[f"resize_{name}_{l}"] = prev
    out[f"resize_{name}_odd"] = cv2.resize(img, (w * 3 // 5, h * 5 // 7), interpolation=cv2.INTER_LINEAR)
    out[f"border_{name}"] = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    out[f"blur_{name}"] = cv2.GaussianBlur(img, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
    for th in (7, 20):
        det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True,
                                             type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        kps = det.detect(img, None)
        out[f"fast_{name}_{th}"] = np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in kps],
                                            np.int32).reshape(-1, 3)
y = np.concatenate([rng.normal(0, 5000, 4000), rng.integers(-3, 4, 200), [0, 0, 1, -1, 7, -7]]).astype(np.float32)
x = np.concatenate([rng.normal(0, 5000, 4000), rng.integers(-3, 4, 200), [0, 5, 0, 0, 7, 7]]).astype(np.float32)
out["atan2_y"], out["atan2_x"] = y, x
out["atan2"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
np.savez_compressed(os.path.join(HERE, "cv2_kat.npz"), **out)
print("wrote", os.path.join(HERE, "cv2_kat.npz"), {k: v.shape for k, v in out.items() if k.startswith("fast")})
