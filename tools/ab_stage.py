#!/usr/bin/env python3
"""A/B timing of kernel builds: for every liborbx build given (ORBX_LIB override, one fresh process each), the per-stage device
times and the whole-step time of one extraction of B synthetic frames resident on the device, plus a hash of the results so that
a build that changes any output byte shows up.
usage: python tools/ab_stage.py [--config 0] [--batch 512] [--reps 20] lib1.so lib2.so ...   (no library: the in-tree build)"""
import argparse, hashlib, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHAPES = {0: (640, 480, 1000, 512), 1: (752, 480, 1200, 384), 2: (1241, 376, 2000, 256), 4: (1920, 1080, 4000, 64)}


def child(args):
    sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
    import numpy as np, torch, orbx
    from orbx import synth
    W, H, NF, B0 = SHAPES[args.config]
    B = args.batch or B0
    ex = orbx.ORBextractor(NF, 1.2, 8, 20, 7, max_batch=B)
    frames = np.stack([synth.rects(W, H, 1000 + i) for i in range(min(B, 64))])
    pitch = (W + 31) // 32 * 32
    slab = np.zeros((B, H, pitch), np.uint8)
    for b in range(B):
        slab[b, :, :W] = frames[b % len(frames)]
    dev = torch.from_numpy(slab).cuda()
    run = lambda: ex.extract_device(dev.data_ptr(), B, W, H, pitch, pitch * H)
    for _ in range(3):
        run()
    ex.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st = torch.cuda.ExternalStream(ex.stream())
    best = 1e9
    for _ in range(3):
        with torch.cuda.stream(st):
            e0.record()
            for _ in range(args.reps):
                run()
            e1.record()
        ex.sync()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / args.reps)
    ex.set_profiling(True)
    acc = {}
    for _ in range(5):
        run()
        for k, v in ex.stage_ms().items():
            acc[k] = acc.get(k, 0.0) + v / 5
    ex.set_profiling(False)
    run()
    res = ex.fetch()
    h = hashlib.sha256()
    for k, d in res:
        h.update(k.tobytes()); h.update(d.tobytes())
    print(json.dumps({"lib": os.path.basename(orbx.LIB_PATH), "ms_per_step": round(best, 4), "frames_per_s": round(B / best * 1e3),
                      "stage_ms": {k: round(v, 4) for k, v in acc.items()}, "results_sha": h.hexdigest()[:12]}))


ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=0)
ap.add_argument("--batch", type=int, default=0)
ap.add_argument("--reps", type=int, default=20)
ap.add_argument("--child", action="store_true")
ap.add_argument("libs", nargs="*")
args = ap.parse_args()
if args.child:
    child(args)
else:
    for lib in args.libs or [""]:
        env = dict(os.environ)
        if lib:
            env["ORBX_LIB"] = os.path.abspath(lib)
        r = subprocess.run([sys.executable, __file__, "--child", "--config", str(args.config), "--batch", str(args.batch), "--reps", str(args.reps)],
                           env=env, capture_output=True, text=True)
        print(r.stdout.strip() or ("FAILED %s: %s" % (lib, r.stderr.strip()[-400:])), flush=True)
