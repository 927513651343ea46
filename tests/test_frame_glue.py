"""Frame glue around the extractor: grey conversion (Tracking.cc:280-305), UndistortKeyPoints
(Frame.cc:419-449), ComputeImageBounds (:451-479), ComputeStereoFromRGBD (:658-679).

The reference delegates the arithmetic to OpenCV (cvtColor, undistortPoints), which it does not vendor: tests/golden/cvtcolor_s9000.npz
and undistort_s9000.npz hold cv2 4.13.0 outputs (tools/gen_golden_glue.py).  CPU leg: oracle == cv2 goldens, bit for bit.
GPU leg: C ABI == goldens and == oracle, bit for bit; colour extraction == grey extraction of the converted image."""
import os

import numpy as np
import pytest

import oracle_py as O
from orbx import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CODES = [("img3", True, "rgb2gray"), ("img3", False, "bgr2gray"), ("img4", True, "rgba2gray"), ("img4", False, "bgra2gray")]


def _cameras():
    g = np.load(os.path.join(GOLD, "undistort_s9000.npz"))
    for n in g["names"]:
        K = g["K_" + n]
        yield str(n), (K[0, 0], K[1, 1], K[0, 2], K[1, 2]), g["D_" + n].reshape(-1), g["pts_" + n], g["und_" + n]


def _color_frame(w, h, seed, cn):
    chans = [synth.rects(w, h, seed + c) for c in range(3)]
    if cn == 4:
        chans.append(np.full((h, w), 255, np.uint8))
    return np.ascontiguousarray(np.stack(chans, 2))


def _keypoints(n, w, h, seed):
    rng = np.random.default_rng(seed)
    k = np.zeros(n, O.KP_DTYPE)
    k["x"], k["y"] = rng.uniform(0, w - 1, n), rng.uniform(0, h - 1, n)
    k["size"], k["angle"], k["response"], k["octave"], k["class_id"] = 31, rng.uniform(0, 360, n), rng.uniform(1, 99, n), rng.integers(0, 8, n), -1
    return k


# ------------------------------------------------------------------------------------------------ CPU: oracle vs cv2 goldens
def test_oracle_cvt_gray_matches_cv2_golden():
    g = np.load(os.path.join(GOLD, "cvtcolor_s9000.npz"))
    assert str(g["cv2_version"]).startswith("4.")
    for src, rgb, key in CODES:
        assert np.array_equal(O.cvt_gray(g[src], rgb), g[key]), key


def test_oracle_undistort_matches_cv2_golden():
    n = 0
    for name, K4, D, pts, und in _cameras():
        got = O.undistort_points(pts, K4, D)
        assert got.tobytes() == und.tobytes(), name
        assert np.abs(und - pts).max() > 1.0          # the distortion moves points by pixels in every case
        n += 1
    assert n == 4


def test_oracle_frame_glue_host_logic():
    name, K4, D, pts, und = next(_cameras())
    k = _keypoints(400, 640, 480, 3)
    k["x"], k["y"] = pts[:, 0], pts[:, 1]
    ku = O.undistort_keypoints(k, K4, D)
    assert np.array_equal(ku["x"], und[:, 0]) and np.array_equal(ku["y"], und[:, 1])
    for f in ("size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(ku[f], k[f])
    assert O.undistort_keypoints(k, K4, np.zeros(4, np.float32)).tobytes() == k.tobytes()          # k1 == 0: mvKeysUn = mvKeys
    b = O.image_bounds(640, 480, K4, D)
    assert b.tolist() == [min(und[0, 0], und[2, 0]), max(und[1, 0], und[3, 0]), min(und[0, 1], und[1, 1]), max(und[2, 1], und[3, 1])]
    assert O.image_bounds(640, 480, K4, np.zeros(5, np.float32)).tolist() == [0.0, 640.0, 0.0, 480.0]
    rng = np.random.default_rng(4)
    depth = np.where(rng.random((480, 640)) < 0.3, 0.0, rng.uniform(0.4, 8.0, (480, 640))).astype(np.float32)
    kk = _keypoints(500, 640, 480, 5)
    kku = O.undistort_keypoints(kk, K4, D)
    uR, dep = O.stereo_from_rgbd(kk, kku, depth, 40.0)
    d = depth[kk["y"].astype(np.int32), kk["x"].astype(np.int32)]
    assert np.array_equal(dep, np.where(d > 0, d, np.float32(-1)))
    with np.errstate(divide="ignore"):
        assert np.array_equal(uR, np.where(d > 0, kku["x"] - np.float32(40.0) / d, np.float32(-1)).astype(np.float32))
    assert (dep > 0).sum() > 100 and (dep < 0).sum() > 50


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_cuda_undistort_matches_cv2_golden_and_oracle():
    import orbx
    m = orbx.ORBmatcher()
    for name, K4, D, pts, und in _cameras():
        assert m.UndistortPoints(pts, K4, D).tobytes() == und.tobytes(), name
        big = np.random.default_rng(7).uniform(-50, 800, (20000, 2)).astype(np.float32)
        assert m.UndistortPoints(big, K4, D).tobytes() == O.undistort_points(big, K4, D).tobytes(), name
        k = _keypoints(len(pts), 640, 480, 3)
        k["x"], k["y"] = pts[:, 0], pts[:, 1]
        assert m.UndistortKeyPoints(k, K4, D).tobytes() == O.undistort_keypoints(k, K4, D).tobytes()
        assert m.UndistortKeyPoints(k, K4, np.zeros(4, np.float32)).tobytes() == k.tobytes()
        cols, rows = (752, 480) if name == "euroc" else (640, 480)
        assert m.ComputeImageBounds(cols, rows, K4, D).tobytes() == O.image_bounds(cols, rows, K4, D).tobytes()
        assert m.ComputeImageBounds(cols, rows, K4, np.zeros(5, np.float32)).tolist() == [0.0, float(cols), 0.0, float(rows)]


@pytest.mark.gpu
def test_cuda_stereo_from_rgbd():
    import orbx
    m = orbx.ORBmatcher()
    name, K4, D, _, _ = next(_cameras())
    rng = np.random.default_rng(4)
    depth = np.where(rng.random((480, 640)) < 0.3, 0.0, rng.uniform(0.4, 8.0, (480, 640))).astype(np.float32)
    k = _keypoints(1500, 640, 480, 5)
    ku = m.UndistortKeyPoints(k, K4, D)
    uR, dep = m.ComputeStereoFromRGBD(k, ku, depth, 40.0)
    ouR, odep = O.stereo_from_rgbd(k, ku, depth, 40.0)
    assert uR.tobytes() == ouR.tobytes() and dep.tobytes() == odep.tobytes()
    padded = np.zeros((480, 648), np.float32)                  # a depth image with a row stride
    padded[:, :640] = depth
    uR2, dep2 = m.ComputeStereoFromRGBD(k, ku, padded[:, :640], 40.0)
    assert uR2.tobytes() == ouR.tobytes() and dep2.tobytes() == odep.tobytes()


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,cn,rgb", [(640, 480, 3, True), (640, 480, 4, False), (321, 243, 3, False), (322, 241, 4, True), (323, 240, 3, True)])
def test_cuda_colour_extraction_equals_grey_extraction(w, h, cn, rgb):
    import orbx
    img = _color_frame(w, h, 100 + w + cn, cn)
    grey = O.cvt_gray(img, rgb)
    ex = orbx.ORBextractor(700, 1.2, 8, 20, 7)
    (kps, desc), = ex.extract_color_batch([img], rgb)
    assert np.array_equal(ex.pyramid_level(0), grey)                                   # the grey image itself
    okps, odesc = O.OracleExtractor(700, 1.2, 8, 20, 7)(grey)
    assert kps.tobytes() == okps.tobytes() and np.array_equal(desc, odesc) and len(kps) > 200
    assert ex.last_launches() == 12


@pytest.mark.gpu
def test_cuda_colour_golden_and_device_batch():
    import torch
    import orbx
    g = np.load(os.path.join(GOLD, "cvtcolor_s9000.npz"))
    frames = [_color_frame(640, 480, 300 + i, 3) for i in range(3)]
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=3)
    host = ex.extract_color_batch(frames, True)
    dev = torch.from_numpy(np.stack(frames)).cuda()
    ex2 = orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=3)
    ex2.extract_color_device(dev.data_ptr(), 3, 640, 480, 640 * 3, 640 * 480 * 3, 3, True)
    res = ex2.fetch()
    oe = O.OracleExtractor(1000, 1.2, 8, 20, 7)
    for b in range(3):
        ok, od = oe(O.cvt_gray(frames[b], True))
        for kps, desc in (host[b], res[b]):
            assert kps.tobytes() == ok.tobytes() and np.array_equal(desc, od)
    # the golden images are smaller than one FAST cell grid allows; check the conversion alone through the pyramid accessor
    for src, rgb, key in CODES:
        big = np.tile(g[src], (8, 6, 1))[:280, :360]
        ex3 = orbx.ORBextractor(300, 1.2, 8, 20, 7)
        ex3.extract_color_batch([big], rgb)
        assert np.array_equal(ex3.pyramid_level(0), np.tile(g[key], (8, 6))[:280, :360]), key
    with pytest.raises(orbx.OrbxError):
        ex.extract_color_device(dev.data_ptr(), 3, 640, 480, 640 * 3, 640 * 480 * 3, 2, True)
