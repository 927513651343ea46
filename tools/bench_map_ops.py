#!/usr/bin/env python3
"""Throughput of the map-side operators through the C ABI with HOST arrays in and out
(that is how their callers hold the data), next to the CPU oracle on one host thread: Fuse / SearchBySim3 (one-kernel window +
gate + argmin), SearchByProjection(Frame, KF), SearchForTriangulation, ComputeDistinctiveDescriptors, ORBVocabulary::transform,
UndistortKeyPoints and the colour ingestion.  Prints one JSON object per operator.  usage: python tools/bench_map_ops.py"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import orbx
import oracle_py as O          # CPU baseline + parity check
import scenes


def timed(fn, reps):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        r = fn()
    return (time.perf_counter() - t0) / reps, r


def line(name, unit, work, t_gpu, t_cpu, **kw):
    print(json.dumps(dict(op=name, unit=unit, gpu=work / t_gpu, cpu_1_thread=work / t_cpu, gpu_ms=t_gpu * 1e3, cpu_ms=t_cpu * 1e3, **kw)), flush=True)


def tracking_searches():
    """The searches Tracking makes on every frame, named by BASELINE.json's north_star: SearchForInitialization (mono
    initialisation, 2 * nFeatures extractor, windowSize 100), SearchByProjection(Frame, local MapPoints),
    SearchByProjection(CurrentFrame, LastFrame), SearchByBoW(KF, Frame) and SearchByBoW(KF, KF)."""
    from orbx import synth
    ex = orbx.ORBextractor(2000, 1.2, 8, 20, 7)
    img1 = synth.rects(640, 480, 1000)
    rng = np.random.default_rng(1000)
    img2 = np.clip(np.roll(img1, 3, axis=1).astype(np.int16) + rng.integers(-3, 4, img1.shape), 0, 255).astype(np.uint8)
    k1, d1 = ex(img1)
    k2, d2 = ex(img2)
    bounds = (0.0, 640.0, 0.0, 480.0)
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    mi = orbx.ORBmatcher(0.9, True)
    tg, (n, m12, _) = timed(lambda: mi.SearchForInitialization(k1, d1, k2, d2, bounds, prev, 100), 30)
    tc, (on, om12, _) = timed(lambda: O.search_for_initialization(k1, d1, k2, d2, bounds, prev, 100, 0.9, True), 5)
    assert n == on and np.array_equal(m12, om12)
    line("SearchForInitialization", "calls/s", 1, tg, tc, features=(len(k1), len(k2)), matches=int(n))
    # the same with both frames resident on the device (orbm_frame; built once per frame, here from host arrays)
    tb, (F1, F2) = timed(lambda: (orbx.DeviceFrame(mi, k1, d1, bounds), orbx.DeviceFrame(mi, k2, d2, bounds)), 20)
    tg, (n, m12, _) = timed(lambda: mi.SearchForInitializationDev(F1, F2, prev, 100), 50)
    assert n == on and np.array_equal(m12, om12)
    line("SearchForInitialization, device frames", "calls/s", 1, tg, tc, features=(len(k1), len(k2)), matches=int(n), frame_build_ms=tb * 1e3 / 2)

    s = scenes.projection_scene(1000)
    a = (s["kps"], s["desc"], s["u_right"], s["occupied"], scenes.BOUNDS, scenes.SF, s["proj_xy"], s["proj_xr"], s["level"])
    mp = orbx.ORBmatcher(0.8, True)
    tg, (n, fm) = timed(lambda: mp.SearchByProjection(*a, s["view_cos"], s["mp_desc"], s["mp_has_obs"], 3.0), 50)
    tc, (on, ofm) = timed(lambda: O.search_by_projection(*a, s["view_cos"], s["mp_desc"], s["mp_has_obs"], 3.0, 0.8), 10)
    assert n == on and np.array_equal(fm, ofm)
    line("SearchByProjection(Frame, MapPoints)", "calls/s", 1, tg, tc, points=len(s["level"]), features=len(s["kps"]), matches=int(n))
    tb, F = timed(lambda: orbx.DeviceFrame(mp, s["kps"], s["desc"], scenes.BOUNDS), 20)
    tg, (n, fm) = timed(lambda: mp.SearchByProjectionDev(F, s["u_right"], s["occupied"], scenes.SF, s["proj_xy"], s["proj_xr"], s["level"], s["view_cos"],
                                                         s["mp_desc"], s["mp_has_obs"], 3.0), 100)
    assert n == on and np.array_equal(fm, ofm)
    line("SearchByProjection(Frame, MapPoints), device frame", "calls/s", 1, tg, tc, points=len(s["level"]), features=len(s["kps"]), matches=int(n),
         frame_build_ms=tb * 1e3)
    mf = orbx.ORBmatcher(0.9, True)
    tg, (n, fm) = timed(lambda: mf.SearchByProjectionFrame(*a, s["angle"], s["mp_desc"], s["mp_has_obs"], 15.0, 0), 50)
    tc, (on, ofm) = timed(lambda: O.search_by_projection_frame(*a, s["angle"], s["mp_desc"], s["mp_has_obs"], 15.0, 0, True), 10)
    assert n == on and np.array_equal(fm, ofm)
    line("SearchByProjection(CurrentFrame, LastFrame)", "calls/s", 1, tg, tc, points=len(s["level"]), features=len(s["kps"]), matches=int(n))
    tg, (n, fm) = timed(lambda: mf.SearchByProjectionFrameDev(F, s["u_right"], s["occupied"], scenes.SF, s["proj_xy"], s["proj_xr"], s["level"], s["angle"],
                                                              s["mp_desc"], s["mp_has_obs"], 15.0, 0), 100)
    assert n == on and np.array_equal(fm, ofm)
    line("SearchByProjection(CurrentFrame, LastFrame), device frame", "calls/s", 1, tg, tc, points=len(s["level"]), features=len(s["kps"]), matches=int(n))

    b = scenes.bow_scene(2000)
    mb = orbx.ORBmatcher(0.7, True)
    tg, (n, mt) = timed(lambda: mb.SearchByBoW(b["k1"], b["d1"], b["valid1"], b["fv1"], b["k2"], b["d2"], b["fv2"], None), 50)
    tc, (on, omt) = timed(lambda: O.search_by_bow(b["k1"], b["d1"], b["valid1"], b["fv1"], b["k2"], b["d2"], b["fv2"], None, 0.7, True), 10)
    assert n == on and np.array_equal(mt, omt)
    line("SearchByBoW(KF, Frame)", "calls/s", 1, tg, tc, features=(len(b["k1"]), len(b["k2"])), matches=int(n))
    mb2 = orbx.ORBmatcher(0.8, True)
    tg, (n, mt) = timed(lambda: mb2.SearchByBoW(b["k1"], b["d1"], b["valid1"], b["fv1"], b["k2"], b["d2"], b["fv2"], b["valid2"]), 50)
    tc, (on, omt) = timed(lambda: O.search_by_bow(b["k1"], b["d1"], b["valid1"], b["fv1"], b["k2"], b["d2"], b["fv2"], b["valid2"], 0.8, True), 10)
    assert n == on and np.array_equal(mt, omt)
    line("SearchByBoW(KF, KF)", "calls/s", 1, tg, tc, features=(len(b["k1"]), len(b["k2"])), matches=int(n))


def main():
    tracking_searches()
    m = orbx.ORBmatcher(0.7, True)
    # ---- Fuse: 8 key frames' worth of points against one key frame (LocalMapping::SearchInNeighbors fuses ~20 x 1000 points)
    s = scenes.fuse_scene(1000)
    rep = 20
    big = {k: (np.concatenate([s[k]] * rep) if k in ("proj_xy", "proj_ur", "level", "mp_desc") else s[k]) for k in s}
    args = (big["kps"], big["desc"], big["u_right"], scenes.BOUNDS, scenes.SF, scenes.INV_SIGMA2, big["proj_xy"], big["proj_ur"], big["level"], big["mp_desc"])
    tg, (n, bi) = timed(lambda: m.Fuse(*args, 3.0, True), 20)
    tc, (on, obi) = timed(lambda: O.fuse(*args, 3.0, True), 3)
    assert n == on and np.array_equal(bi, obi)
    line("Fuse(KF, MapPoints)", "points/s", len(big["level"]), tg, tc, points=len(big["level"]), features=len(big["kps"]), fused=int(n))
    # ---- SearchBySim3
    s3 = scenes.sim3_scene(1000)
    a3 = (s3["k1"], s3["d1"], s3["k2"], s3["d2"], scenes.BOUNDS, scenes.SF, s3["proj12_xy"], s3["level12"], s3["mp1_desc"], s3["proj21_xy"], s3["level21"], s3["mp2_desc"])
    tg, (n, mt) = timed(lambda: m.SearchBySim3(*a3, 7.5), 50)
    tc, (on, omt) = timed(lambda: O.search_by_sim3(*a3, 7.5), 10)
    assert n == on and np.array_equal(mt, omt)
    line("SearchBySim3", "calls/s", 1, tg, tc, features=(len(s3["k1"]), len(s3["k2"])), found=int(n))
    # ---- SearchByProjection(Frame, KF)
    sp = scenes.projection_scene(1000)
    ap = (sp["kps"], sp["desc"], sp["occupied"], scenes.BOUNDS, scenes.SF, sp["proj_xy"], sp["level"], sp["angle"], sp["mp_desc"])
    tg, (n, fm) = timed(lambda: m.SearchByProjectionKF(*ap, 10.0, 100), 50)
    tc, (on, ofm) = timed(lambda: O.search_by_projection_kf(*ap, 10.0, 100, True), 10)
    assert n == on and np.array_equal(fm, ofm)
    line("SearchByProjection(Frame, KF)", "calls/s", 1, tg, tc, points=len(sp["level"]), matches=int(n))
    # ---- SearchForTriangulation
    st = scenes.triangulation_scene(1000)
    at = (st["k1"], st["d1"], st["has_mp1"], st["u_right1"], st["fv1"], st["k2"], st["d2"], st["has_mp2"], st["u_right2"], st["fv2"], st["F12"],
          st["epipole"], scenes.SF, scenes.SIGMA2)
    tg, (n, mt) = timed(lambda: m.SearchForTriangulation(*at, False), 50)
    tc, (on, omt) = timed(lambda: O.search_for_triangulation(*at, False, True), 10)
    assert n == on and np.array_equal(mt, omt)
    line("SearchForTriangulation", "calls/s", 1, tg, tc, features=(len(st["k1"]), len(st["k2"])), matches=int(n))
    # ---- ComputeDistinctiveDescriptors: 20000 MapPoints with 2..30 observations
    rng = np.random.default_rng(1)
    sizes = rng.integers(2, 31, 20000)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    desc = rng.integers(0, 256, (int(off[-1]), 32)).astype(np.uint8)
    tg, best = timed(lambda: m.ComputeDistinctiveDescriptors(desc, off), 10)
    tc, obest = timed(lambda: O.distinctive_descriptors(desc, off), 2)
    assert np.array_equal(best, obest)
    line("ComputeDistinctiveDescriptors", "MapPoints/s", len(sizes), tg, tc, observations=int(off[-1]))
    # ---- vocabulary transform: ORBvoc geometry (k = 10, L = 6 would be 1.1 M nodes; L = 5 here: 111 k nodes), 2000 descriptors / frame
    (k1, d1), (k2, d2) = scenes.frames(1000, 2000)
    k, L, parent, vdesc, weight = scenes.synthetic_vocabulary(3, 10, 5, shallow_leaf_prob=0.02, desc_pool=d1)
    v = orbx.ORBVocabulary(k, L, parent, vdesc, weight)
    ov = O.OracleVocabulary(k, L, parent, vdesc, weight)
    feats = np.concatenate([d1, d2])[:2000]
    tg, ((bid, bval), fv) = timed(lambda: v.transform(feats, 4), 50)
    tc, ((obid, obval), ofv) = timed(lambda: ov.transform(feats, 4), 5)
    assert np.array_equal(bid, obid) and bval.tobytes() == obval.tobytes() and all(np.array_equal(a, b) for a, b in zip(fv, ofv))
    line("ORBVocabulary::transform", "descriptors/s", len(feats), tg, tc, nodes=len(parent), k=k, L=L)
    # ---- UndistortKeyPoints
    K4, D = (517.306408, 516.469215, 318.643040, 255.313989), np.array([0.262383, -0.953104, -0.005358, 0.002628, 1.163314], np.float32)
    tg, ku = timed(lambda: m.UndistortKeyPoints(k1, K4, D), 100)
    tc, oku = timed(lambda: O.undistort_keypoints(k1, K4, D), 100)
    assert ku.tobytes() == oku.tobytes()
    line("UndistortKeyPoints", "keypoints/s", len(k1), tg, tc, keypoints=len(k1))
    # ---- colour ingestion: 128 RGB frames resident on the device, extraction included vs grey input
    B = 128
    col = np.stack([np.stack([orbx.synth.rects(640, 480, 10 * i + c) for c in range(3)], 2) for i in range(8)])
    dev = torch.from_numpy(np.ascontiguousarray(col[np.arange(B) % 8])).cuda()
    grey = torch.from_numpy(np.ascontiguousarray(np.stack([O.cvt_gray(col[i], True) for i in range(8)])[np.arange(B) % 8])).cuda()
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B)
    st_ = torch.cuda.ExternalStream(ex.stream())

    def dev_ms(fn, reps=10):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st_)
        for _ in range(reps):
            fn()
        e1.record(st_)
        e1.synchronize()
        return e0.elapsed_time(e1) / reps

    t_col = dev_ms(lambda: ex.extract_color_device(dev.data_ptr(), B, 640, 480, 640 * 3, 640 * 480 * 3, 3, True))
    r_col = ex.fetch()
    t_gry = dev_ms(lambda: ex.extract_device(grey.data_ptr(), B, 640, 480, 640, 640 * 480))
    r_gry = ex.fetch()
    assert all(a[0].tobytes() == b[0].tobytes() and np.array_equal(a[1], b[1]) for a, b in zip(r_col, r_gry))
    print(json.dumps(dict(op="colour ingestion (RGB -> grey -> extraction, device resident)", frames_per_s_rgb=B / (t_col / 1e3), frames_per_s_grey=B / (t_gry / 1e3),
                          grey_conversion_us_per_frame=(t_col - t_gry) * 1e3 / B, hbm_gbs_of_conversion=B * 640 * 480 * 4 / ((t_col - t_gry) / 1e3) / 1e9)), flush=True)


if __name__ == "__main__":
    main()
