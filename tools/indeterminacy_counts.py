#!/usr/bin/env python3
"""Counts the reference's two compile/allocator-dependent behaviours on the bench frames, with the
CPU oracle: (a) (frame, level) cases where the heap-address tie-break of the largest-first quadtree phase
(ORBextractor.cc:684, tie_mode 1 = real list nodes under glibc malloc) selects different keypoints from the
creation-order definition (tie_mode 0); (b) descriptor bits that differ between the FMA-contracted rotation g++ -O3
-march=native emits for ORBextractor.cc:119-120 (fma_mode 1) and the uncontracted definition (fma_mode 0)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200")); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import oracle_py as O
from orbx import synth
N = int(sys.argv[1]) if len(sys.argv) > 1 else 64
a, b, c = O.OracleExtractor(1000, 1.2, 8, 20, 7), O.OracleExtractor(1000, 1.2, 8, 20, 7, tie_mode=1), O.OracleExtractor(1000, 1.2, 8, 20, 7, fma_mode=1)
tie_cases = tie_frames = 0
bits = rows = nkp = 0
for i in range(N):
    img = synth.rects(640, 480, 1000 + i)
    k0, d0 = a(img); k1, d1 = b(img); k2, d2 = c(img)
    lv = 0
    for l in range(8):
        m0, m1 = k0[k0["octave"] == l], k1[k1["octave"] == l]
        if len(m0) != len(m1) or m0.tobytes() != m1.tobytes():
            lv += 1
    tie_cases += lv; tie_frames += lv > 0
    assert k0.tobytes() == k2.tobytes()
    x = np.unpackbits(d0 ^ d2, axis=1)
    bits += int(x.sum()); rows += int(x.any(axis=1).sum()); nkp += len(k0)
print("frames %d: pointer tie-break differs in %d (frame, level) cases of %d (%d frames affected)" % (N, tie_cases, 8 * N, tie_frames))
print("FMA-contracted vs uncontracted rotation: %d descriptor bits in %d of %d descriptors (%.4f %% of bits)" % (bits, rows, nkp, 100.0 * bits / (256.0 * nkp)))
