#!/usr/bin/env python3
"""Randomised parity sweep: random image shapes (odd sizes, narrow and tall), extractor parameters and frame kinds, CUDA
extractor (single calls and small batches, host and 4-byte-aligned device input) against the CPU oracle, bit for bit.
usage: python tools/fuzz_parity.py [seconds] [seed]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import torch
import orbx
from orbx import synth
import oracle_py as O          # checker

budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)

# ---- brute-force matcher (a fifth of the budget): random set sizes (0, 1, ragged, several segments), clustered descriptors so
# that equal distances and equal best / second occur; the same matcher handle throughout (its ticket counters must return to 0)
t_end = time.time() + budget / 5
m = orbx.ORBmatcher(0.9, True)
nmatch = 0
while time.time() < t_end:
    nq, nt = int(rng.choice([0, 1, 31, 129, 777, 2000, 4100])), int(rng.choice([0, 1, 2, 33, 64, 65, 500, 3000, 9000]))
    base = rng.integers(0, 256, (8, 32)).astype(np.uint8)
    def cluster(n):
        d = base[rng.integers(0, 8, n)].copy()
        flips = rng.integers(0, 256, (n, 6))
        on = rng.random((n, 6)) < 0.5
        for j in range(6):
            d[np.arange(n), flips[:, j] >> 3] ^= (on[:, j] * (1 << (flips[:, j] & 7))).astype(np.uint8)
        return d
    q, t = cluster(nq), cluster(nt)
    bi, bd, sd = m.match_bruteforce(q, t)
    obi, obd, osd = O.match_bruteforce(q, t)
    assert np.array_equal(bi, obi) and np.array_equal(bd, obd) and np.array_equal(sd, osd), ("bruteforce", nq, nt)
    nmatch += 1
print("fuzz ok (matcher): %d brute-force problems index- and distance-equal to the oracle" % nmatch)
t_end = time.time() + budget
ncases = nframes = 0
while time.time() < t_end:
    w, h = int(rng.integers(96, 900)), int(rng.integers(96, 700))
    nf = int(rng.choice([50, 300, 1000, 2000, 3000]))
    sf = float(rng.choice([1.1, 1.2, 1.3, 1.5]))
    nl = int(rng.integers(1, 9))
    ini = int(rng.integers(8, 40))
    mn = int(rng.integers(2, ini + 1))
    kind = rng.choice(["rects", "texture", "lowcontrast"])
    try:
        ex = orbx.ORBextractor(nf, sf, nl, ini, mn, max_batch=3)
        oe = O.OracleExtractor(nf, sf, nl, ini, mn)
        imgs = [getattr(synth, kind)(w, h, int(rng.integers(1, 1 << 30))) for _ in range(3)]
        ref = []
        for im in imgs:
            try:
                ref.append(oe(im))
            except Exception:
                ref = None                       # shapes the reference itself cannot process (level smaller than a cell)
                break
        if ref is None:
            try:
                ex(imgs[0])
                raise AssertionError("CUDA path accepted a shape the oracle rejects: %dx%d nl=%d sf=%g" % (w, h, nl, sf))
            except orbx.OrbxError:
                continue
        for rep in range(2):                     # the second round replays the captured graph
            for im, (ok, od) in zip(imgs, ref):
                k, d = ex(im)
                assert k.tobytes() == ok.tobytes() and np.array_equal(d, od), ("host", w, h, nf, sf, nl, ini, mn, kind)
        pitch = (w + 3) & ~3
        dev = torch.zeros((3, h, pitch), dtype=torch.uint8, device="cuda")
        for i, im in enumerate(imgs):
            dev[i, :, :w] = torch.from_numpy(im).cuda()
        torch.cuda.synchronize()
        for form in (0, 1, 2, 3):
            ex.set_octree_form(form)
            ex.extract_device(dev.data_ptr(), 3, w, h, pitch, pitch * h)
            res = ex.fetch()
            for (k, d), (ok, od) in zip(res, ref):
                assert k.tobytes() == ok.tobytes() and np.array_equal(d, od), ("device", form, w, h, nf, sf, nl, ini, mn, kind)
        ncases += 1
        nframes += 3
    except orbx.OrbxError as e:
        if "quadtree needs" in str(e) or "smaller than one" in str(e) or "height > 2*width" in str(e):
            continue
        raise
print("fuzz ok: %d configurations, %d distinct frames, every result bit-equal to the oracle" % (ncases, nframes))
