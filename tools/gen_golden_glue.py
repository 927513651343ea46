#!/usr/bin/env python3
"""Golden vectors for the Frame glue around the extractor, produced with cv2 4.13.0 in the build
container and committed under tests/golden/: the reference calls these OpenCV functions (cvtColor at Tracking.cc:280-305,
undistortPoints at Frame.cc:419-449) and pins no OpenCV version, so the wheel here is the anchor.  Inputs are stored (they are tiny)."""
import os
import numpy as np
import cv2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = os.path.join(ROOT, "tests", "golden")
rng = np.random.default_rng(9000)

# ---- cvtColor -> GRAY: random bytes plus the extreme triples
img3 = rng.integers(0, 256, (37, 61, 3), dtype=np.uint8)
img3[0, :8] = [[0, 0, 0], [255, 255, 255], [255, 0, 0], [0, 255, 0], [0, 0, 255], [255, 255, 0], [0, 255, 255], [255, 0, 255]]
img4 = rng.integers(0, 256, (37, 61, 4), dtype=np.uint8)
np.savez_compressed(os.path.join(out, "cvtcolor_s9000.npz"), cv2_version=cv2.__version__, img3=img3, img4=img4,
                    rgb2gray=cv2.cvtColor(img3, cv2.COLOR_RGB2GRAY), bgr2gray=cv2.cvtColor(img3, cv2.COLOR_BGR2GRAY),
                    rgba2gray=cv2.cvtColor(img4, cv2.COLOR_RGBA2GRAY), bgra2gray=cv2.cvtColor(img4, cv2.COLOR_BGRA2GRAY))

# ---- undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) as in Frame::UndistortKeyPoints / ComputeImageBounds
cases = []
for name, (fx, fy, cx, cy), dist, (w, h) in [
        ("tum1", (517.306408, 516.469215, 318.643040, 255.313989), [0.262383, -0.953104, -0.005358, 0.002628, 1.163314], (640, 480)),
        ("tum2", (520.908620, 521.007327, 325.141442, 249.701764), [0.231222, -0.784899, -0.003257, -0.000105, 0.917205], (640, 480)),
        ("euroc", (458.654, 457.296, 367.215, 248.375), [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05], (752, 480)),
        ("strong", (400.0, 410.0, 300.0, 250.0), [-0.45, 0.35, 0.01, -0.008, -0.12], (640, 480))]:
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    D = np.array(dist, np.float32).reshape(-1, 1)
    pts = np.stack([rng.uniform(0, w, 400), rng.uniform(0, h, 400)], 1).astype(np.float32)
    pts[:4] = [[0, 0], [w, 0], [0, h], [w, h]]                     # the corners ComputeImageBounds undistorts
    und = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, D, None, K).reshape(-1, 2)
    cases.append((name, K, D, pts, und))
np.savez_compressed(os.path.join(out, "undistort_s9000.npz"), cv2_version=cv2.__version__, names=[c[0] for c in cases],
                    **{"K_%s" % c[0]: c[1] for c in cases}, **{"D_%s" % c[0]: c[2] for c in cases},
                    **{"pts_%s" % c[0]: c[3] for c in cases}, **{"und_%s" % c[0]: c[4] for c in cases})
print("written", cv2.__version__)
