#!/usr/bin/env python3
"""One process, every GPU: orbx_pool_extract_batch (the library's multi-GPU driver) on 1, 2, 4, ... GPUs of the box, host frames in
page-locked memory -> host results.  Weak scaling: 512 frames per GPU per call.  Also checks that the first 96 frames come back
byte-identical for every device count.  usage: tools/pool_scale.py [--config 0|4] [--lanes 4] [--chunk 32] [--steps 6]"""
import argparse, ctypes as C, hashlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
import numpy as np, torch, orbx
from orbx import synth
ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=0)
ap.add_argument("--lanes", type=int, default=4)
ap.add_argument("--chunk", type=int, default=0)
ap.add_argument("--steps", type=int, default=6)
a = ap.parse_args()
W, H, NF, per = {0: (640, 480, 1000, 512), 4: (1920, 1080, 4000, 64)}[a.config]
chunk = a.chunk or (32 if a.config == 0 else 8)
L = orbx.lib()
present = torch.cuda.device_count()
P32 = (W + 31) & ~31
frame = (P32 * H + 255) & ~255
distinct = np.stack([synth.rects(W, H, 1000 + i) for i in range(16)])
out, ref = [], None
G = 1
while G <= present:
    B = per * G
    pin = torch.zeros((B, frame), dtype=torch.uint8).pin_memory()
    arr = pin.numpy()
    for i in range(B):
        arr[i, :P32 * H].reshape(H, P32)[:, :W] = distinct[i % 16]
    pool = orbx.ExtractorPool(NF, 1.2, 8, 20, 7, devices=list(range(G)), chunk=chunk, lanes=a.lanes)
    cap = pool.plan(W, H)
    pk = torch.zeros((B, cap, 28), dtype=torch.uint8).pin_memory()
    pd = torch.zeros((B, cap, 32), dtype=torch.uint8).pin_memory()
    n = np.zeros(B, np.int32)
    ptrs = (C.c_void_p * B)(*[pin.data_ptr() + i * frame for i in range(B)])

    def call():
        rc = L.orbx_pool_extract_batch(pool._h, ptrs, B, W, H, P32, C.c_void_p(pk.data_ptr()), C.c_void_p(pd.data_ptr()), cap, n.ctypes.data_as(C.c_void_p))
        assert rc == 0, L.orbx_last_error()
    call(); call()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        call()
    dt = (time.perf_counter() - t0) / a.steps
    h = hashlib.sha256()
    for b in range(min(96, per)):
        h.update(pk[b].numpy()[:n[b]].tobytes()); h.update(pd[b].numpy()[:n[b]].tobytes())
    ref = ref or h.hexdigest()
    out.append({"gpus": G, "frames_per_call": B, "frames_per_s": B / dt, "ms_per_call": dt * 1e3, "h2d_GBps": B * ((H - 1) * P32 + W) / dt / 1e9,
                "first_frames_identical_to_1_gpu": h.hexdigest() == ref})
    print(out[-1], file=sys.stderr, flush=True)
    pool.close()
    del pin, pk, pd
    G *= 2
print(json.dumps({"tool": "pool_scale", "config": a.config, "shape": [W, H, NF], "lanes": a.lanes, "chunk": chunk, "host_threads": os.cpu_count(), "runs": out}))
