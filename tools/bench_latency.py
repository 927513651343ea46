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

# ---- the C++ classes themselves (include/ORBextractor.h, include/ORBmatcher.h), used the way the reference uses them: tools/latency_cpp.cpp
import subprocess, tempfile
tmp = tempfile.mkdtemp()
exe = os.path.join(tmp, "latency_cpp")
libdir = os.path.dirname(orbx.LIB_PATH)
subprocess.check_call(["g++", "-std=c++11", "-O2", "-I", os.path.join(ROOT, "include"), "-o", exe, os.path.join(ROOT, "tools", "latency_cpp.cpp"),
                       "-L", libdir, "-lorbx", "-Wl,-rpath," + libdir])
for (w, h, nf) in [(640, 480, 1000), (640, 480, 2000)]:
    a = synth.rects(w, h, 1000)
    rng = np.random.default_rng(1000)
    b = np.clip(np.roll(a, 3, axis=1).astype(np.int16) + rng.integers(-3, 4, a.shape), 0, 255).astype(np.uint8)
    open(os.path.join(tmp, "a.raw"), "wb").write(a.tobytes())
    open(os.path.join(tmp, "b.raw"), "wb").write(b.tobytes())
    out = subprocess.run([exe, os.path.join(tmp, "a.raw"), os.path.join(tmp, "b.raw"), str(w), str(h), str(nf)], capture_output=True, text=True)
    if out.returncode != 0:
        raise SystemExit("latency_cpp failed: " + out.stderr[-2000:])
    for ln in out.stdout.splitlines():
        d = json.loads(ln)
        d["config"] = "C++ classes, %dx%d nf=%d" % (w, h, nf)
        print(json.dumps(d), flush=True)
