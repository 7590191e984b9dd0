"""Generates tests/golden/cv_golden.npz from cv2 (OpenCV 4.13.0 wheel, IPP on) -- the third-party
arithmetic the reference calls on this path:
  cv::getStructuringElement(MORPH_ELLIPSE)    collision_checking.cpp:106-107
  cv::dilate                                   collision_checking.cpp:223
  cv::distanceTransform(DIST_L2, DIST_MASK_5)  a_star.cpp:429   (IPP float and plain 16.16 variants)
Run here (container with cv2); the .npz is committed, this script is its provenance."""
import os

import cv2
import numpy as np

out = {}
out["cv_version"] = np.array(cv2.__version__)
for r in (1, 2, 3, 4, 7, 10, 11, 12, 16):
    d = 2 * r + 1
    out["se_%d" % r] = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (d, d), (r, r))

rng = np.random.default_rng(1234)
for i, (dim, r, p) in enumerate([(96, 10, 0.004), (131, 3, 0.02), (200, 10, 0.002), (64, 16, 0.01)]):
    img = np.zeros((dim, dim), np.uint8)
    img[rng.random((dim, dim)) < p] = 1
    img[rng.random((dim, dim)) < p] = 2
    d = 2 * r + 1
    se = cv2.getStructuringElement(cv2.MORPH_ELLIPSE, (d, d), (r, r))
    out["dil_in_%d" % i] = img
    out["dil_r_%d" % i] = np.array(r)
    out["dil_out_%d" % i] = cv2.dilate(img, se)

for i, (h, w, p) in enumerate([(200, 200, 0.01), (200, 200, 0.0005), (120, 90, 0.03), (200, 200, 0.2)]):
    src = np.ones((h, w), np.uint8)
    src[rng.random((h, w)) < p] = 0
    if i == 3:  # rectangles, like a pooled obstacle grid
        src[:] = 1
        for _ in range(12):
            x, y = rng.integers(0, w - 20), rng.integers(0, h - 20)
            src[y:y + rng.integers(2, 20), x:x + rng.integers(2, 20)] = 0
    out["dt_src_%d" % i] = src
    cv2.ipp.setUseIPP(True)
    out["dt_ipp_%d" % i] = cv2.distanceTransform(src, cv2.DIST_L2, cv2.DIST_MASK_5)
    cv2.ipp.setUseIPP(False)
    out["dt_fix_%d" % i] = cv2.distanceTransform(src, cv2.DIST_L2, cv2.DIST_MASK_5)
    cv2.ipp.setUseIPP(True)
out["ipp_used"] = np.array(bool(cv2.ipp.useIPP()))
np.savez_compressed(os.path.join(os.path.dirname(os.path.abspath(__file__)), "cv_golden.npz"), **out)
print("written", len(out), "arrays; ipp", cv2.ipp.useIPP(), cv2.ipp.getIppVersion())
