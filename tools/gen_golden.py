#!/usr/bin/env python3
"""Generate the golden fixtures under tests/golden/ (run in the build container; outputs are committed).

The reference ships no golden vectors, so these are produced by the CPU oracle AFTER it has been
checked, frame by frame, against the independent Python+cv2 4.13 orchestration (oracle/cv2_cross_oracle.py): a fixture
is written only when the two agree byte for byte.  Inputs are not stored: they are regenerated from the seeded
generators in orbx.synth (the fixture records kind, shape and seed)."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
import oracle_py as O
import cv2_cross_oracle as X
from orbx import synth

CASES = [("rects", 640, 480, 1000, 1000), ("texture", 640, 480, 1001, 1000), ("lowcontrast", 640, 480, 1002, 1000),
         ("rects", 752, 480, 2000, 1200), ("rects", 320, 240, 7, 500)]
out = os.path.join(ROOT, "tests", "golden")
for kind, w, h, seed, nf in CASES:
    img = synth.frame(kind, w, h, seed)
    oe = O.OracleExtractor(nf, 1.2, 8, 20, 7)
    kps, desc = oe(img)
    ckps, cdesc = X.Cv2Extractor(nf, 1.2, 8, 20, 7)(img)
    ref = np.array(ckps, dtype=np.float64).reshape(-1, 7)
    assert len(kps) == len(ckps) and np.array_equal(desc, cdesc)
    for i, f in enumerate(["x", "y", "size", "angle", "response", "octave", "class_id"]):
        assert np.array_equal(kps[f].astype(np.float64), ref[:, i]), f
    cand = [oe.level_candidates(l) for l in range(8)]
    np.savez_compressed(os.path.join(out, "extract_%s_%dx%d_s%d_nf%d.npz" % (kind, w, h, seed, nf)),
                        kind=kind, w=w, h=h, seed=seed, nf=nf, kps=kps, desc=desc,
                        level_counts=np.array([oe.level_nkeys(l) for l in range(8)]),
                        cand_counts=np.array([len(c) for c in cand]),
                        cand_l0=cand[0], cand_l7=cand[7],
                        level7=oe.level_padded(7), blurred7=oe.level_blurred(7))
    print(kind, w, h, len(kps))
# matcher: config 4 at N = 2000 (best index, best and second distance per query)
A, B = synth.descriptor_sets(2000, 6000)
bi, bd, sd = O.match_bruteforce(A, B)
np.savez_compressed(os.path.join(out, "match_bruteforce_2000_s6000.npz"), N=2000, seed=6000, best_idx=bi, best_dist=bd, second_dist=sd)
print("ok")
