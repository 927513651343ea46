#!/usr/bin/env python3
"""Per-stage device times of one extraction configuration: tools/stage_times.py W H nfeatures B [steps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
import numpy as np
import torch
import orbx
from orbx import synth
W, H, nf, B = (int(v) for v in sys.argv[1:5])
steps = int(sys.argv[5]) if len(sys.argv) > 5 else 5
pitch = (W + 15) & ~15
distinct = [synth.rects(W, H, 5000 + i) for i in range(4)]
dev = torch.zeros((B, H, pitch), dtype=torch.uint8, device="cuda")
for i in range(B):
    dev[i, :, :W] = torch.from_numpy(distinct[i % 4]).cuda()
ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7, max_batch=B)
if os.environ.get("ORBX_FORM"):
    ex.set_octree_form(int(os.environ["ORBX_FORM"]))      # 1 / 2 / 3 = 128 / 256 / 512 threads
for _ in range(2):
    ex.extract_device(dev.data_ptr(), B, W, H, pitch, pitch * H)
ex.sync()
ex.set_profiling(True)
acc = {}
for _ in range(steps):
    ex.extract_device(dev.data_ptr(), B, W, H, pitch, pitch * H)
    ex.sync()
    for k, v in ex.stage_ms().items():
        acc[k] = acc.get(k, 0.0) + v / steps
tot = sum(acc.values())
print("%dx%d nf=%d B=%d: %.0f frames/s" % (W, H, nf, B, B / tot * 1e3), {k: round(v, 3) for k, v in acc.items()})
