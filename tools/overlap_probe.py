#!/usr/bin/env python3
"""Does splitting a batch over several handles (streams) raise device-resident throughput?  tools/overlap_probe.py [B]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
import numpy as np, torch, orbx
from orbx import synth
B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
W, H = 640, 480
distinct = np.stack([synth.rects(W, H, 1000 + i) for i in range(32)])
host = np.ascontiguousarray(np.tile(distinct, (B // 32 + 1, 1, 1))[:B])
dev = torch.from_numpy(host).cuda()
for parts in (1, 2, 4):
    sub = B // parts
    exs = [orbx.ORBextractor(1000, 1.2, 8, 20, 7, max_batch=sub) for _ in range(parts)]
    def step():
        for i, ex in enumerate(exs):
            ex.extract_device(dev.data_ptr() + i * sub * W * H, sub, W, H, W, W * H)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    n = 20
    for _ in range(n):
        step()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / n
    print("B=%d as %d x %d: %.3f ms per step, %.0f frames/s" % (B, parts, sub, dt * 1e3, B / dt), flush=True)
    del exs
