#!/usr/bin/env python3
"""CPU fuzz of the oracle's extractor restatement against the UNMODIFIED reference src/ORBextractor.cc (oracle/_ref/liborb_ref*.so, node
allocator 1 = the deterministic tie-break): random image kinds, sizes (odd ones included), feature counts, scale factors, level counts and
FAST thresholds.  Keypoints (all seven fields) and descriptors must be identical, for the reference's flags (fma_mode 1) and for
-ffp-contract=off (fma_mode 0).  usage: python tools/fuzz_reference_extractor.py [runs] [seed]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py as O, ref_py as R
from orbx import synth

runs, seed = int(sys.argv[1]) if len(sys.argv) > 1 else 100, int(sys.argv[2]) if len(sys.argv) > 2 else 1
rng = np.random.default_rng(seed)
bad = 0
for r in range(runs):
    w, h = int(rng.integers(80, 900)), int(rng.integers(80, 700))
    if rng.random() < 0.3:
        w, h = [(640, 480), (752, 480), (1241, 376), (320, 240), (1280, 720)][int(rng.integers(0, 5))]
    nf = int(rng.choice([50, 200, 500, 1000, 1200, 2000, 4000]))
    sf = float(rng.choice([1.2, 1.2, 1.1, 1.3, 1.44, 2.0]))
    nlev = int(rng.integers(1, 11))
    while min(w, h) / sf ** (nlev - 1) < 45:      # the reference needs every level to hold the 19-px border and a FAST cell
        nlev -= 1
    # DistributeOctTree divides by nIni = round(width / height) of the bordered level (ORBextractor.cc:543-545): 0, and a crash in the
    # reference, where a level is more than twice as high as wide; the product rejects such shapes
    if any((round(w / sf ** l) - 26) / max(round(h / sf ** l) - 26, 1) < 0.52 for l in range(nlev)):
        w, h = h, w
    ini, mn = [(20, 7), (20, 7), (30, 10), (12, 5), (7, 7), (40, 3)][int(rng.integers(0, 6))]
    kind = ["rects", "texture", "lowcontrast", "flat"][int(rng.choice(4, p=[0.45, 0.35, 0.15, 0.05]))]
    s = int(rng.integers(0, 1 << 20))
    img = synth.frame(kind, w, h, s)
    res = []
    for fma, variant in ((1, "fma"), (0, "nofma")):
        ko, do = O.OracleExtractor(nf, sf, nlev, ini, mn, tie_mode=0, fma_mode=fma)(img)
        kr, dr = R.RefExtractor(nf, sf, nlev, ini, mn, allocator=1, variant=variant)(img)
        res.append(ko.tobytes() == kr.tobytes() and do.tobytes() == dr.tobytes())
    ok = all(res)
    bad += not ok
    print("%3d %-11s %4dx%-4d nf %4d sf %.2f levels %2d th %2d/%-2d seed %7d: %4d keypoints: %s" % (r, kind, w, h, nf, sf, nlev, ini, mn, s, len(ko), "identical" if ok else "DIFFERENT %r" % res), flush=True)
print("%d runs, %d with a difference" % (runs, bad))
sys.exit(1 if bad else 0)
