#!/usr/bin/env python3
"""CPU fuzz of the vocabulary transform -- the oracle's restatement and the drop-in include/ORBVocabulary.h (over the C ABI served by the
oracle) -- against the vendored DBoW2 (oracle/_ref/libframe_ref.so): random branching factors, depths, scorings and weightings; BowVector
ids and doubles, FeatureVector, and score(v1, v2) bit for bit.  L2 scoring is reported apart: BowVector::normalize sums v*v, which the
reference's flags contract into fma.  usage: python tools/fuzz_reference_vocabulary.py [runs] [seed]"""
import os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for d in ("oracle", "orb_slam2v2-1_b200", "tests"):
    sys.path.insert(0, os.path.join(ROOT, d))
import oracle_py as O, ref_frame_py as RF, scenes

runs, seed = int(sys.argv[1]) if len(sys.argv) > 1 else 40, int(sys.argv[2]) if len(sys.argv) > 2 else 4
rng = np.random.default_rng(seed)
tmp = tempfile.mkdtemp()
assert RF.dropin_available(cpu=True)
(k1, d1), (k2, d2) = scenes.frames(1000)
feats = np.concatenate([d1, d2, rng.integers(0, 256, (400, 32)).astype(np.uint8)])
bad = l2 = 0
for r in range(runs):
    k = int(rng.integers(2, 21)); L = int(rng.integers(1, 5 if k > 8 else 7))
    while k ** L > 60000:
        L -= 1
    sc, wt, s = int(rng.integers(0, 6)), int(rng.integers(0, 4)), int(rng.integers(0, 1000))
    voc = scenes.synthetic_vocabulary(s, k, L, shallow_leaf_prob=0.0, desc_pool=d1)
    path = os.path.join(tmp, "v.txt")
    scenes.write_vocabulary_text(path, *voc, sc, wt)
    with open(path, "rb+") as f:                       # no trailing newline: see tests/test_reference_frame.py
        f.seek(-1, os.SEEK_END); f.truncate()
    rv, ov, pv = RF.RefVocabulary(path), O.OracleVocabulary(*voc, sc, wt), RF.RefVocabulary(path, libpath=RF.DROPIN_CPU_PATH)
    ids = vals = True
    for lu in (0, 1, 2, 4, 7):
        (b, v), fv = rv.transform(feats, lu); (ob, ovv), ofv = ov.transform(feats, lu); (pb, pvv), pfv = pv.transform(feats, lu)
        ids = ids and np.array_equal(b, ob) and np.array_equal(b, pb) and all(np.array_equal(x, y) and np.array_equal(x, z) for x, y, z in zip(fv, ofv, pfv))
        vals = vals and v.tobytes() == ovv.tobytes() == pvv.tobytes()
        close = np.allclose(v, ovv, rtol=4e-16, atol=0) and np.allclose(v, pvv, rtol=4e-16, atol=0)
    va, vb = rv.transform(d1, 2)[0], rv.transform(d2, 2)[0]
    score = np.float64(rv.score(va, vb)).tobytes() == np.float64(pv.score(va, vb)).tobytes()
    ok = ids and score and (vals or (sc == 1 and close))
    l2 += ok and not vals
    bad += not ok
    print("%3d k %2d L %d scoring %d weighting %d words %5d: %s" % (r, k, L, sc, wt, rv.size(), "identical" if ok and vals else ("identical but for the last bit of the L2-normalised values" if ok else "DIFFERENT")), flush=True)
print("%d runs, %d with a difference, %d L2 cases off by the last bit of the norm" % (runs, bad, l2))
sys.exit(1 if bad else 0)
