#!/usr/bin/env python3
"""Fuzz of the oracle's Frame-level restatements against Frames built by the UNMODIFIED reference src/Frame.cc (oracle/_ref/libframe_ref.so):
random shapes, feature counts, baselines and cameras.  Stereo: mvuRight / mvDepth of Frame::ComputeStereoMatches; RGB-D: mvKeysUn, image
bounds, mvuRight / mvDepth of ComputeStereoFromRGBD; grid: GetFeaturesInArea.  CPU only.
usage: python tools/fuzz_reference_frame.py [runs] [seed]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as O, ref_frame_py as RF
from orbx import synth

runs, seed = int(sys.argv[1]) if len(sys.argv) > 1 else 30, int(sys.argv[2]) if len(sys.argv) > 2 else 1
rng = np.random.default_rng(seed)
bad = 0
for r in range(runs):
    w, h = [(640, 480), (752, 480), (512, 384), (400, 300), (800, 600), (1241, 376), (1280, 720)][int(rng.integers(0, 7))]
    nf = int(rng.choice([300, 500, 1000, 1200, 2000, 3000]))
    nlev, sf = [(8, 1.2), (8, 1.2), (6, 1.3), (10, 1.15), (4, 1.5)][int(rng.integers(0, 5))]
    fx = float(rng.uniform(250, 900)); bf = float(rng.uniform(5, 400)); s = int(rng.integers(0, 1 << 20))
    K4 = [fx, fx * float(rng.uniform(0.97, 1.03)), w * 0.5 + float(rng.uniform(-8, 8)), h * 0.5 + float(rng.uniform(-8, 8))]
    kind = r % 3
    try:
        if kind == 0:
            L, R = synth.stereo_rects(w, h, s)[:2]
            F = RF.RefFrame("stereo", L, K4, np.zeros(4), bf=bf, im_right=R, nfeatures=nf, scaleFactor=sf, nlevels=nlev)
            n, uR, dep = O.stereo_matches(F.keys, F.desc, F.keys_right, F.desc_right, F.pyramid(0), F.pyramid(1), F.sf, F.isf, bf, float(np.float32(bf) / F.mb))
            ok = uR.tobytes() == F.u_right.tobytes() and dep.tobytes() == F.depth_out.tobytes()
            what = "stereo %d matches" % n
        elif kind == 1:
            dist = [float(rng.uniform(-0.35, 0.35)), float(rng.uniform(-1, 1)), float(rng.uniform(-0.01, 0.01)), float(rng.uniform(-0.01, 0.01))] + \
                   ([float(rng.uniform(-1.2, 1.2))] if rng.random() < 0.5 else [])
            depth = rng.uniform(0.3, 12.0, (h, w)).astype(np.float32)
            depth[rng.random((h, w)) < 0.25] = 0.0
            F = RF.RefFrame("rgbd", synth.rects(w, h, s), K4, dist, bf=bf, depth=depth, nfeatures=nf, scaleFactor=sf, nlevels=nlev)
            K4f = [np.float32(v) for v in K4]
            uR, dep = O.stereo_from_rgbd(F.keys, F.keys_un, depth, bf)
            ok = (O.undistort_keypoints(F.keys, K4f, dist).tobytes() == F.keys_un.tobytes() and O.image_bounds(w, h, K4f, dist).tobytes() == F.bounds.tobytes()
                  and uR.tobytes() == F.u_right.tobytes() and dep.tobytes() == F.depth_out.tobytes())
            what = "rgbd %d coefficients" % len(dist)
        else:
            dist = [float(rng.uniform(-0.3, 0.3)), float(rng.uniform(-0.8, 0.8)), 0.001, -0.002] if rng.random() < 0.5 else [0.0] * 4
            F = RF.RefFrame("mono", synth.texture(w, h, s) if rng.random() < 0.5 else synth.rects(w, h, s), K4, dist, bf=bf, nfeatures=nf, scaleFactor=sf, nlevels=nlev)
            G = O.lib().orbo_grid_build(O._p(F.keys_un), F.N, *[float(b) for b in F.bounds])
            buf = np.zeros(max(F.N, 1), np.int32)
            ok, nq = True, 0
            for q in range(150):
                x, y = float(rng.uniform(-40, w + 40)), float(rng.uniform(-40, h + 40))
                rad = float(rng.choice([2.0, 7.5, 15.0, 40.0, 120.0]))
                lo, hi = [(-1, -1), (0, -1), (1, -1), (0, 2), (1, 3), (-1, 2)][int(rng.integers(0, 6))]
                n = O.lib().orbo_grid_query(G, O._p(F.keys_un), x, y, rad, lo, hi, O._p(buf), len(buf))
                ok = ok and buf[:n].tolist() == F.features_in_area(x, y, rad, lo, hi).tolist()
                nq += n > 0
            O.lib().orbo_grid_destroy(G)
            what = "grid %d non-empty queries" % nq
    except Exception as e:      # noqa
        ok, what = False, "exception %r" % (e,)
    bad += not ok
    print("%3d %4dx%-4d nf %4d levels %2d x %.2f: %s: %s" % (r, w, h, nf, nlev, sf, what, "identical" if ok else "DIFFERENT"), flush=True)
print("%d runs, %d with a difference" % (runs, bad))
sys.exit(1 if bad else 0)
