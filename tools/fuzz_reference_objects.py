#!/usr/bin/env python3
"""Fuzz of tests/native/reference_objects_test.cpp on the CPU: random scene kinds, sizes, feature counts and shifts; every run compares all
eleven searches of the drop-in ORBmatcher (over the C ABI served by the CPU oracle) with the reference's on the reference's real objects.
usage: python tools/fuzz_reference_objects.py <exe> [runs] [seed]      (exe: built as in tests/test_reference_matcher.py::real_cpu_exe)"""
import os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
from orbx import synth

exe, runs, seed = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40, int(sys.argv[3]) if len(sys.argv) > 3 else 1
rng = np.random.default_rng(seed)
tmp = tempfile.mkdtemp()
bad = 0
for r in range(runs):
    w, h = [(640, 480), (752, 480), (512, 384), (400, 300), (800, 600), (1241, 376)][int(rng.integers(0, 6))]
    nf = int(rng.choice([300, 500, 1000, 1200, 2000, 3000]))
    kind = ["rects", "texture"][int(rng.integers(0, 2))]
    dx, dy = int(rng.integers(-7, 8)), int(rng.integers(-5, 6))
    s = int(rng.integers(0, 1 << 20))
    a = getattr(synth, kind)(w, h, s)
    b = np.clip(np.roll(a, (dy, dx), axis=(0, 1)).astype(np.int16) + np.random.default_rng(s).integers(-3, 4, a.shape), 0, 255).astype(np.uint8)
    a.tofile(os.path.join(tmp, "a.raw")); b.tofile(os.path.join(tmp, "b.raw"))
    p = subprocess.run([exe, os.path.join(tmp, "a.raw"), os.path.join(tmp, "b.raw"), str(w), str(h), str(nf), str(dx), str(dy)], capture_output=True, text=True)
    # "too few ..." checks guard against vacuous comparisons on the fixed test scenes; a random scene may legitimately be sparse
    fails = [l for l in p.stderr.splitlines() if l.startswith("FAIL") and "too few" not in l]
    ok = not fails and p.returncode in (0, 4) and "error" not in p.stderr
    bad += not ok
    print("%3d %s %4dx%-4d nf %4d shift (%+d,%+d) seed %7d: %s" % (r, kind, w, h, nf, dx, dy, s, "identical" if ok else "DIFFERENT " + "; ".join(fails)[:300] + p.stderr[-200:]), flush=True)
print("%d runs, %d with a difference" % (runs, bad))
sys.exit(1 if bad else 0)
