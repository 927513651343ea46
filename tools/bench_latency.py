#!/usr/bin/env python3
"""Latency of ONE ORBextractor::operator() call -- how the reference is used (one frame per call, Frame.cc:247-253) -- through
the C ABI with host arrays in and out, next to the CPU oracle on one thread.  usage: python tools/bench_latency.py"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import orbx
from orbx import synth
import oracle_py as O          # CPU baseline + parity check

for (w, h, nf) in [(640, 480, 1000), (752, 480, 1200), (1241, 376, 2000), (1920, 1080, 4000)]:
    imgs = [synth.rects(w, h, 7000 + i) for i in range(8)]
    ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    if os.environ.get("ORBX_FORM"):
        ex.set_octree_form(int(os.environ["ORBX_FORM"]))      # 1 / 2 / 3 = 128 / 256 / 512 threads
    oe = O.OracleExtractor(nf, 1.2, 8, 20, 7)
    k, d = ex(imgs[0])
    ok, od = oe(imgs[0])
    assert k.tobytes() == ok.tobytes() and np.array_equal(d, od)
    for i in range(20):
        ex(imgs[i % 8])
    ts = []
    for i in range(200):
        t0 = time.perf_counter()
        ex(imgs[i % 8])
        ts.append(time.perf_counter() - t0)
    tc = []
    for i in range(8):
        t0 = time.perf_counter()
        oe(imgs[i % 8])
        tc.append(time.perf_counter() - t0)
    print(json.dumps({"config": "%dx%d nf=%d, one frame per call" % (w, h, nf), "gpu_ms_median": float(np.median(ts)) * 1e3,
                      "gpu_ms_p99": float(np.percentile(ts, 99)) * 1e3, "cpu_1_thread_ms_median": float(np.median(tc)) * 1e3,
                      "launches": ex.last_launches(), "keypoints": int(len(k))}), flush=True)
