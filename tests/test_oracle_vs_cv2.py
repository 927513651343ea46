"""Pins the CPU oracle (oracle/liborb_oracle.so) to OpenCV 4.13 primitive semantics and to the independent
Python+cv2 orchestration (oracle/cv2_cross_oracle.py).  The reference ships no tests or golden vectors
(SURVEY.md section 4), so this is what anchors the oracle."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
import oracle_py as O
from orbx import synth

SHAPES = [(640, 480), (752, 480), (1241, 376)]


def _rand(w, h, seed):
    return np.random.default_rng(seed).integers(0, 256, (h, w), dtype=np.uint8)


@pytest.mark.parametrize("w,h", SHAPES + [(1920, 1080)])
def test_resize_chain_matches_cv2(w, h):
    ex = O.OracleExtractor()
    isf = ex.scale_tables()[1]
    cur = _rand(w, h, 1)
    for l in range(1, 8):
        dw = int(np.rint(np.float32(w) * isf[l])); dh = int(np.rint(np.float32(h) * isf[l]))
        ref = cv2.resize(cur, (dw, dh), interpolation=cv2.INTER_LINEAR)
        got = O.resize_linear(cur, dw, dh)
        assert np.array_equal(ref, got), (l, int((ref != got).sum()))
        cur = ref


@pytest.mark.parametrize("sw,sh,dw,dh", [(64, 48, 53, 40), (33, 31, 32, 30), (100, 80, 99, 41), (40, 40, 57, 61), (9, 9, 7, 7)])
def test_resize_odd_ratios(sw, sh, dw, dh):
    src = _rand(sw, sh, 3)
    assert np.array_equal(cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR), O.resize_linear(src, dw, dh))


def test_border_matches_cv2():
    src = _rand(57, 41, 2)
    assert np.array_equal(cv2.copyMakeBorder(src, 19, 19, 19, 19, cv2.BORDER_REFLECT_101), O.border_reflect101(src, 19))


@pytest.mark.parametrize("w,h", [(300, 200), (179, 134), (31, 17), (640, 480)])
def test_gauss_matches_cv2(w, h):
    for seed, img in enumerate([_rand(w, h, 5), synth.rects(w, h, 9) if w >= 100 else _rand(w, h, 6)]):
        ref = cv2.GaussianBlur(img.copy(), (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(ref, O.gauss7(img)), seed


def test_fast_atan2_matches_cv2():
    rng = np.random.default_rng(7)
    ys = rng.integers(-2_900_000, 2_900_000, 50000)
    xs = rng.integers(-2_900_000, 2_900_000, 50000)
    ys[:200] = rng.integers(-3, 4, 200); xs[:200] = rng.integers(-3, 4, 200)
    bad = 0
    for y, x in zip(ys, xs):
        a = np.float32(cv2.fastAtan2(float(y), float(x)))
        b = np.float32(O.fast_atan2(float(y), float(x)))
        bad += a.tobytes() != b.tobytes()
    assert bad == 0
    assert O.fast_atan2(0.0, 0.0) == 0.0


def _cv_fast(roi, t):
    det = cv2.FastFeatureDetector_create(int(t), True)
    return np.array([(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in det.detect(roi)], dtype=np.int32).reshape(-1, 3)


def test_fast_cells_match_cv2():
    rng = np.random.default_rng(11)
    base = [synth.rects(640, 480, 1000), synth.texture(640, 480, 3), _rand(640, 480, 8),
            cv2.GaussianBlur(_rand(640, 480, 9), (5, 5), 1.2)]
    n = 0
    for trial in range(160):
        img = base[trial % 4]
        w, h = int(rng.integers(7, 44)), int(rng.integers(7, 44))
        x0, y0 = int(rng.integers(0, 640 - w)), int(rng.integers(0, 480 - h))
        roi = img[y0:y0 + h, x0:x0 + w]       # a strided sub-view, like rowRange().colRange()
        for t in (20, 7):
            ref = _cv_fast(roi, t)
            got = O.fast(roi, t, True)
            assert np.array_equal(ref, got), (trial, t, len(ref), len(got))
            n += len(ref)
    assert n > 500


def test_fast_tiny_roi_is_empty():
    img = _rand(64, 64, 4)
    assert len(O.fast(img[:6, :40], 7)) == 0 and len(_cv_fast(img[:6, :40], 7)) == 0
    assert len(O.fast(img[:40, :6], 7)) == 0


@pytest.mark.parametrize("kind,w,h,seed,nf", [("rects", 640, 480, 1000, 1000), ("texture", 640, 480, 1001, 1000),
                                             ("lowcontrast", 640, 480, 1002, 1000), ("rects", 752, 480, 2000, 1200)])
def test_full_frame_matches_cv2_orchestration(kind, w, h, seed, nf):
    import cv2_cross_oracle as X
    img = synth.frame(kind, w, h, seed)
    ex = O.OracleExtractor(nf, 1.2, 8, 20, 7)
    kps, desc = ex(img)
    cx = X.Cv2Extractor(nf, 1.2, 8, 20, 7)
    ckps, cdesc = cx(img)
    assert list(ex.features_per_level()) == cx.per_level
    assert list(ex.umax()) == cx.umax
    for l in range(8):
        assert np.array_equal(ex.level_padded(l), cx.padded[l]), l
        got = ex.level_candidates(l)
        ref = np.array([c[:3] for c in cx.candidates[l]], dtype=np.int32).reshape(-1, 3)
        assert np.array_equal(got, ref), l
    assert len(kps) == len(ckps)
    ref = np.array(ckps, dtype=np.float64)
    for i, f in enumerate(["x", "y", "size", "angle", "response", "octave", "class_id"]):
        assert np.array_equal(kps[f].astype(np.float64), ref[:, i]), f
    assert np.array_equal(desc, cdesc)


def test_flat_frame_has_no_keypoints():
    kps, desc = O.OracleExtractor()(synth.flat(640, 480))
    assert len(kps) == 0 and desc.shape == (0, 32)
