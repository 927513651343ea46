#!/usr/bin/env python3
"""One warm-up extraction plus one profiled extraction of B synthetic 640x480 frames (for ncu):
ncu --set full --import-source on --clock-control none --launch-skip 22 --launch-count 22 -o gpurun_out/x python tools/profile_once.py [B]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
import numpy as np
import torch
import orbx
from orbx import synth
B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
W, H = 640, 480
distinct = np.stack([synth.rects(W, H, 1000 + i) for i in range(32)])
host = np.ascontiguousarray(np.tile(distinct, (B // 32 + 1, 1, 1))[:B])
dev = torch.from_numpy(host).cuda()
ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=B)
for _ in range(2):
    ex.extract_device(dev.data_ptr(), B, W, H, W, W * H)
    ex.sync()
print("ok", ex.last_launches())
