"""MapPoint::ComputeDistinctiveDescriptors (MapPoint.cc:252-312) from the UNMODIFIED reference src/MapPoint.cc
(oracle/_ref/libmappoint_ref.so, oracle/ref_shim/mappoint_ref_api.cpp) against the oracle's restatement, which the GPU suite holds the CUDA
orbm_distinctive_descriptors equal to (tests/test_map_searches.py).  The same reference objects also give UpdateNormalAndDepth, the
distance-invariance getters and PredictScale, which the stand-in MapPoint of oracle/ref_shim/slam/slam_objects.h restates for the matcher
and Frame builds: checked here against the reference's own.  CPU only."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import oracle_py as O

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "..", "oracle", "_ref", "libmappoint_ref.so")


def _lib():
    if not os.path.exists(LIB) and os.path.exists("/root/reference/src/MapPoint.cc"):
        subprocess.check_call(["make", "-C", os.path.join(HERE, "..", "oracle"), "_ref/libmappoint_ref.so"])
    if not os.path.exists(LIB):
        pytest.skip("oracle/_ref/libmappoint_ref.so is not built and /root/reference is not here")
    L = C.CDLL(LIB)
    L.refp_map_point.restype = C.c_int
    L.refp_map_point.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_void_p, C.c_int,
                                 C.c_void_p, C.c_void_p, C.c_void_p]
    return L


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _reference_point(L, pos, desc, octave, centers, bad, sf=1.2, nlevels=8, dists=()):
    desc, octave = np.ascontiguousarray(desc, np.uint8), np.ascontiguousarray(octave, np.int32)
    centers, bad = np.ascontiguousarray(centers, np.float32), np.ascontiguousarray(bad, np.uint8)
    pos, dists = np.ascontiguousarray(pos, np.float32), np.ascontiguousarray(dists, np.float32)
    out_desc, out5, levels = np.full(32, 0xEE, np.uint8), np.zeros(5, np.float32), np.zeros(max(2 * len(dists), 1), np.int32)
    has = L.refp_map_point(_p(pos), len(desc), _p(desc), _p(octave), _p(centers), _p(bad), sf, nlevels, _p(dists), len(dists), _p(out_desc), _p(out5), _p(levels))
    return has, out_desc, out5, levels[:2 * len(dists)].reshape(-1, 2)


def _observations(rng, n, spread):
    base = rng.integers(0, 256, 32).astype(np.uint8)
    d = np.repeat(base[None], n, 0)
    for i in range(n):
        for b in rng.integers(0, 256, int(rng.integers(0, spread + 1))):
            d[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return d


def test_compute_distinctive_descriptors_is_the_reference():
    L = _lib()
    rng = np.random.default_rng(21)
    cases = 0
    for n in [1, 2, 3, 4, 5, 7, 8, 12, 20, 33, 64]:
        for spread in (0, 3, 40, 120):                                     # 0: all observations identical (every median ties)
            d = _observations(rng, n, spread)
            if spread == 3 and n > 3:
                d[n // 2] = d[0]                                            # duplicates inside a varied set
            bad = (rng.random(n) < 0.25).astype(np.uint8) if n > 2 else np.zeros(n, np.uint8)
            has, got, _, _ = _reference_point(L, [0.3, -0.2, 4.0], d, np.zeros(n), rng.normal(0, 0.3, (n, 3)), bad)
            good = d[bad == 0]
            best = O.distinctive_descriptors(good, np.array([0, len(good)], np.int32))[0]
            if len(good) == 0:
                assert best == -1 and has == 0                              # the reference returns before assigning mDescriptor (:282-283)
            else:
                assert has == 1 and got.tobytes() == good[best].tobytes(), (n, spread)
            cases += 1
    assert cases == 44


def test_normal_depth_and_predict_scale_of_the_stand_ins_are_the_reference():
    """UpdateNormalAndDepth (MapPoint.cc:340-383), GetMin/MaxDistanceInvariance (:385-395), PredictScale (:397-429): what the stand-in
    MapPoint of ref_shim/slam/slam_objects.h restates (0.8 f * min, 1.2 f * max, ceil(log(max / d) / logScaleFactor) clamped)."""
    L = _lib()
    rng = np.random.default_rng(22)
    for trial in range(40):
        n = int(rng.integers(1, 9))
        nlevels = int(rng.choice([4, 8, 12]))
        sf = float(rng.choice([1.2, 1.1, 1.44]))
        pos = rng.normal(0, 1, 3).astype(np.float32) + np.float32([0, 0, 5])
        centers = rng.normal(0, 0.5, (n, 3)).astype(np.float32)
        octave = rng.integers(0, nlevels, n).astype(np.int32)
        dists = np.concatenate([rng.uniform(0.05, 60.0, 30), [1e-3, 1e4]]).astype(np.float32)
        has, _, out5, levels = _reference_point(L, pos, _observations(rng, n, 10), octave, centers, np.zeros(n, np.uint8), sf, nlevels, dists)
        assert has == 1
        # distances: ||Pos - Ow_ref|| with cv::norm (double accumulation, one rounding), then float products (:369-379)
        pc = (pos - centers[0]).astype(np.float32)
        dist = np.float32(np.sqrt(np.sum(pc.astype(np.float64) ** 2)))
        sfs = [np.float32(1.0)]
        for _ in range(1, nlevels):
            sfs.append(np.float32(sfs[-1] * np.float32(sf)))
        maxd = np.float32(dist * sfs[octave[0]])
        mind = np.float32(maxd / sfs[nlevels - 1])
        assert out5[3] == np.float32(np.float32(0.8) * mind) and out5[4] == np.float32(np.float32(1.2) * maxd)
        logsf = np.float32(np.log(np.float32(sf)))                          # Frame.cc:187 / KeyFrame: log(mfScaleFactor) as float
        for q, dq in enumerate(dists):
            ratio = np.float32(maxd / dq)
            lv = int(np.ceil(np.float64(np.float32(np.log(np.float64(ratio)))) / np.float64(logsf)))
            lv = min(max(lv, 0), nlevels - 1)
            # the level is an integer step function of a float quotient: allow the restated arithmetic to sit on the other side of a
            # step only where the quotient is within one ulp of an integer
            quo = np.log(np.float64(ratio)) / np.float64(logsf)
            near_step = abs(quo - round(quo)) < 1e-5
            assert levels[q, 0] == levels[q, 1]                             # KeyFrame and Frame variants agree
            assert levels[q, 0] == lv or near_step, (trial, q, levels[q], lv, quo)
        # the normal: mean of the unit viewing rays (:359-380)
        rays = pos[None].astype(np.float64) - centers.astype(np.float64)
        normal = (rays / np.linalg.norm(rays, axis=1, keepdims=True)).mean(0)
        assert np.allclose(out5[:3], normal, atol=2e-6)
