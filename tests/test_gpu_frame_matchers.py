"""Parity of the frame-level searches (grid candidates, SearchForInitialization, ComputeStereoMatches) with the oracle."""
import numpy as np
import pytest

import oracle_py as O
import orbx
from orbx import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def m():
    return orbx.ORBmatcher(0.9, True)


def _extract_pair(w, h, nf, seed):
    L, R = synth.stereo_rects(w, h, seed)
    exL, exR = orbx.ORBextractor(nf, 1.2, 8, 20, 7), orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    kl, dl = exL(L)
    kr, dr = exR(R)
    return exL, exR, kl, dl, kr, dr


def test_grid_candidates_match_get_features_in_area(m):
    ex = orbx.ORBextractor(2000, 1.2, 8, 20, 7)
    kps, _ = ex(synth.rects(640, 480, 1000))
    bounds = (0.0, 640.0, 0.0, 480.0)
    rng = np.random.default_rng(1)
    q = np.concatenate([np.stack([kps["x"], kps["y"]], 1)[:400] + rng.normal(0, 5, (400, 2)).astype(np.float32),
                        np.array([[-50, -50], [700, 500], [0, 0], [639.9, 479.9], [320, -200], [5000, 240]], np.float32)])
    for (r, lo, hi) in [(100.0, 0, 0), (15.0, -1, -1), (40.0, 2, 5), (7.5, 3, -1), (30.0, 0, 7)]:
        off, idx = m.grid_candidates(kps, bounds, q, r, lo, hi)
        ooff, oidx = O.grid_query_all(kps, q, r, lo, hi, bounds)
        assert np.array_equal(off, ooff) and np.array_equal(idx, oidx), (r, lo, hi)
        assert off[-1] > 0
    # undistorted-style bounds that do not start at 0, per-query radii
    b2 = (-12.5, 655.25, -8.0, 490.5)
    rad = rng.uniform(5, 60, len(q)).astype(np.float32)
    off, idx = m.grid_candidates(kps, b2, q, rad, -1, -1)
    tot = 0
    for i in range(len(q)):
        o, ii = O.grid_query_all(kps, q[i:i + 1], float(rad[i]), -1, -1, b2)
        assert np.array_equal(idx[off[i]:off[i + 1]], ii), i
        tot += len(ii)
    assert tot == off[-1]


@pytest.mark.parametrize("seed,shift", [(1000, 3.0), (1001, 11.0), (1002, 0.0)])
def test_search_for_initialization(m, seed, shift):
    # mono initialisation uses the 2*nFeatures extractor (Tracking.cc:125) and windowSize 100 (Tracking.cc:743)
    ex = orbx.ORBextractor(2000, 1.2, 8, 20, 7)
    img1 = synth.rects(640, 480, seed)
    img2 = np.roll(img1, int(shift), axis=1) if shift else img1.copy()
    rng = np.random.default_rng(seed)
    img2 = np.clip(img2.astype(np.int16) + rng.integers(-3, 4, img2.shape), 0, 255).astype(np.uint8)
    k1, d1 = ex(img1)
    k2, d2 = ex(img2)
    bounds = (0.0, 640.0, 0.0, 480.0)
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    for ratio, ori in [(0.9, True), (0.9, False), (0.7, True)]:
        mm = orbx.ORBmatcher(ratio, ori)
        n, m12, newprev = mm.SearchForInitialization(k1, d1, k2, d2, bounds, prev, 100)
        on, om12, oprev = O.search_for_initialization(k1, d1, k2, d2, bounds, prev, 100, ratio, ori)
        assert n == on and np.array_equal(m12, om12) and newprev.tobytes() == oprev.tobytes()
        assert n > 50
    # second call with the updated vbPrevMatched and a small window, as Tracking does on later frames
    n2, m12b, _ = m.SearchForInitialization(k1, d1, k2, d2, bounds, newprev, 10)
    on2, om12b, _ = O.search_for_initialization(k1, d1, k2, d2, bounds, newprev, 10, 0.9, True)
    assert n2 == on2 and np.array_equal(m12b, om12b)


def test_search_for_initialization_empty(m):
    k = np.zeros(0, orbx.KP_DTYPE)
    d = np.zeros((0, 32), np.uint8)
    n, m12, _ = m.SearchForInitialization(k, d, k, d, (0, 640, 0, 480), np.zeros((0, 2), np.float32), 100)
    assert n == 0 and len(m12) == 0


@pytest.mark.parametrize("w,h,nf,fx,bf,seed", [(752, 480, 1200, 435.2, 47.9, 2000), (1241, 376, 2000, 718.856, 386.1448, 3000),
                                               (752, 480, 1200, 435.2, 47.9, 2001)])
def test_stereo_matches(m, w, h, nf, fx, bf, seed):
    exL, exR, kl, dl, kr, dr = _extract_pair(w, h, nf, seed)
    n, uR, dep = m.ComputeStereoMatches(exL, exR, kl, dl, kr, dr, bf, fx)
    pyrL = [exL.pyramid_level(l) for l in range(8)]
    pyrR = [exR.pyramid_level(l) for l in range(8)]
    on, ouR, odep = O.stereo_matches(kl, dl, kr, dr, pyrL, pyrR, exL.GetScaleFactors(), exL.GetInverseScaleFactors(), bf, fx)
    assert n == on
    assert uR.tobytes() == ouR.tobytes() and dep.tobytes() == odep.tobytes()
    assert n > len(kl) // 10, (n, len(kl))


def test_stereo_no_right_keypoints(m):
    exL, exR, kl, dl, kr, dr = _extract_pair(752, 480, 1200, 2002)
    n, uR, dep = m.ComputeStereoMatches(exL, exR, kl, dl, kr[:0], dr[:0], 47.9, 435.2)
    assert n == 0 and (uR == -1).all() and (dep == -1).all()


@pytest.mark.parametrize("w,h,nf,fx,bf,seed0", [(752, 480, 1200, 435.2, 47.9, 2000), (1241, 376, 2000, 718.856, 386.1448, 3000)])
def test_stereo_matches_batch_device_resident(m, w, h, nf, fx, bf, seed0):
    """orbm_stereo_matches_batch: every pair of a batch, keypoints and pyramids taken from the handles' device results."""
    B = 3
    pairs = [synth.stereo_rects(w, h, seed0 + i) for i in range(B)]
    exL, exR = orbx.ORBextractor(nf, 1.2, 8, 20, 7, max_batch=B), orbx.ORBextractor(nf, 1.2, 8, 20, 7, max_batch=B)
    resL = exL.extract_batch([p[0] for p in pairs])
    resR = exR.extract_batch([p[1] for p in pairs])
    nm, uR, dep = m.ComputeStereoMatchesBatch(exL, exR, bf, fx)
    for b in range(B):
        kl, dl = resL[b]
        kr, dr = resR[b]
        pyrL = [exL.pyramid_level(l, b) for l in range(8)]
        pyrR = [exR.pyramid_level(l, b) for l in range(8)]
        on, ouR, odep = O.stereo_matches(kl, dl, kr, dr, pyrL, pyrR, exL.GetScaleFactors(), exL.GetInverseScaleFactors(), bf, fx)
        assert nm[b] == on, b
        assert uR[b, :len(kl)].tobytes() == ouR.tobytes() and dep[b, :len(kl)].tobytes() == odep.tobytes(), b
        assert on > len(kl) // 10
