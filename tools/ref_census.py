#!/usr/bin/env python3
"""Counters measured on the reference's OWN code (oracle/_ref = the unmodified src/ORBextractor.cc) over the 64 'rects' bench frames
(seeds 1000..1063, 640x480, nfeatures 1000): CPU only, run in the container that holds /root/reference.
  * (frame, level) cases where the stock build (plain malloc: the pair<int, ExtractorNode*> sort of ORBextractor.cc:684 follows the heap's
    history) selects other keypoints than the monotone-allocator definition the oracle and the CUDA path are held to;
  * keypoints both runs keep: any field or descriptor byte differing (expected 0: the difference is the selection only);
  * descriptor bits that differ between the reference compiled with its own flags (FMA-contracted :119-120) and with -ffp-contract=off;
  * oracle restatement vs the reference build on every frame.
usage: python tools/ref_census.py > profiles/r02_ref_census.json"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
import numpy as np
import oracle_py as O, ref_py as R
from orbx import synth
assert R.build(), "oracle/_ref is not built and /root/reference is absent"
P = (1000, 1.2, 8, 20, 7)
stock, mono, nofma, ora = R.RefExtractor(*P, allocator=0), R.RefExtractor(*P, allocator=1), R.RefExtractor(*P, allocator=1, variant="nofma"), O.OracleExtractor(*P, tie_mode=0, fma_mode=1)
levels = differ = frames_differ = shared = shared_bad = kp_total = kp_only_stock = bits = rows = oracle_bad = 0
for i in range(64):
    img = synth.rects(640, 480, 1000 + i)
    (k0, d0), (k1, d1), (k2, d2), (ko, do) = stock(img), mono(img), nofma(img), ora(img)
    oracle_bad += not (ko.tobytes() == k1.tobytes() and np.array_equal(do, d1))
    assert k1.tobytes() == k2.tobytes()
    bits += int(np.unpackbits(d1 ^ d2).sum()); rows += len(d1)
    fd = False
    for l in range(8):
        levels += 1
        a, b = k0[k0["octave"] == l], k1[k1["octave"] == l]
        if a.tobytes() != b.tobytes():
            differ += 1; fd = True
    frames_differ += fd
    pos = {(int(k["octave"]), float(k["x"]), float(k["y"])): j for j, k in enumerate(k0)}
    kp_total += len(k1)
    for j, k in enumerate(k1):
        m = pos.get((int(k["octave"]), float(k["x"]), float(k["y"])))
        if m is None:
            kp_only_stock += 1
        else:
            shared += 1
            shared_bad += (k0[m].tobytes() != k.tobytes()) or not np.array_equal(d0[m], d1[j])
print(json.dumps({"frames": 64, "shape": "640x480 nf=1000 'rects' seeds 1000..1063", "reference": R.lib().orbr_build_info().decode(),
                  "levels": levels, "levels_where_stock_malloc_selects_other_keypoints": differ, "frames_with_any_difference": frames_differ,
                  "keypoints_of_the_definition": kp_total, "of_those_also_kept_by_the_stock_build": shared,
                  "of_those_not_kept_by_the_stock_build": kp_only_stock, "shared_keypoints_with_any_differing_field_or_descriptor_byte": shared_bad,
                  "descriptor_bits_fma_vs_nofma_reference_builds": bits, "descriptor_bits_total": rows * 256,
                  "frames_where_the_oracle_differs_from_the_reference_build": oracle_bad}, indent=1))
