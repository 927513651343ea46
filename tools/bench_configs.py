#!/usr/bin/env python3
"""Throughput of the BASELINE.json configurations other than the headline one (bench.py covers configs[0]):
stereo pairs (752x480 nf=1200, 1241x376 nf=2000: extraction of both images + ComputeStereoMatches) and 1920x1080 nf=4000
frames, device-resident, on one GPU, with the CPU oracle (checker; here the measured baseline) timed beside them.
Prints one JSON object per configuration.  usage: python tools/bench_configs.py [--cpu-seconds S]"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import torch

import orbx
from orbx import synth
import oracle_py as O          # CPU baseline + parity check


def alg_bytes(w, h, nf):
    ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    ex(np.zeros((h, w), np.uint8))
    S = sum(ex.level_size(l)[0] * ex.level_size(l)[1] for l in range(8))
    Spad = sum((ex.level_size(l)[0] + 38) * (ex.level_size(l)[1] + 38) for l in range(8))
    return w * h + Spad + S + 2 * S + min(S, 1369 * nf) + 60 * nf


def cpu_rate(frames, nf, seconds, threads):
    done = [0] * threads
    stop = [0.0]

    def work(t):
        oe = O.OracleExtractor(nf, 1.2, 8, 20, 7)
        i = t
        while time.perf_counter() < stop[0]:
            oe(frames[i % len(frames)])
            done[t] += 1
            i += threads

    O.OracleExtractor(nf, 1.2, 8, 20, 7)(frames[0])
    t0 = time.perf_counter()
    stop[0] = t0 + seconds
    th = [threading.Thread(target=work, args=(t,)) for t in range(threads)]
    [x.start() for x in th]
    [x.join() for x in th]
    return sum(done) / (time.perf_counter() - t0)


def gpu_extract_rate(host, nf, steps=10):
    B, H, W = host.shape
    pitch = (W + 3) & ~3                       # the device entry point wants rows 4-byte aligned
    dev = torch.zeros((B, H, pitch), dtype=torch.uint8, device="cuda")
    dev[:, :, :W] = torch.from_numpy(host).cuda()
    ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7, max_batch=B)
    st = torch.cuda.ExternalStream(ex.stream())
    for _ in range(3):
        ex.extract_device(dev.data_ptr(), B, W, H, pitch, pitch * H)
    ex.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(steps):
        ex.extract_device(dev.data_ptr(), B, W, H, pitch, pitch * H)
    e1.record(st)
    e1.synchronize()
    ms = e0.elapsed_time(e1) / steps
    res = ex.fetch()
    ex._keep = dev                             # the handle's level 0 is this buffer: keep it alive with the handle
    return B / (ms / 1e3), res, ex


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cpu-seconds", type=float, default=6.0)
    args = ap.parse_args()
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("hbm_gbs", 6650.0))
    cores = os.cpu_count() or 1

    # ---- stereo configurations
    for (w, h, nf, seed0, fx, bf, B) in [(752, 480, 1200, 2000, 435.2, 47.9, 192), (1241, 376, 2000, 3000, 718.856, 386.1448, 128)]:
        pairs = [synth.stereo_rects(w, h, seed0 + i) for i in range(8)]
        L = np.ascontiguousarray(np.stack([pairs[i % 8][0] for i in range(B)]))
        R = np.ascontiguousarray(np.stack([pairs[i % 8][1] for i in range(B)]))
        rateL, resL, exL = gpu_extract_rate(L, nf)
        rateR, resR, exR = gpu_extract_rate(R, nf)
        # parity of the first pair against the oracle
        oe = O.OracleExtractor(nf, 1.2, 8, 20, 7)
        okl, odl = oe(L[0])
        assert resL[0][0].tobytes() == okl.tobytes() and np.array_equal(resL[0][1], odl)
        # stereo association per pair on the device pyramids of the two handles (host keypoints in, uRight/depth out)
        m = orbx.ORBmatcher(0.6, True)
        nP = 32
        m.ComputeStereoMatches(exL, exR, resL[0][0], resL[0][1], resR[0][0], resR[0][1], bf, fx, frame=0)      # warm-up: allocates the arena and its page-locked mirror
        t0 = time.perf_counter()
        nm = 0
        for b in range(nP):
            n_, uR, dep = m.ComputeStereoMatches(exL, exR, resL[b][0], resL[b][1], resR[b][0], resR[b][1], bf, fx, frame=b)
            nm += n_
        stereo_s = (time.perf_counter() - t0) / nP
        # the batched, device-resident form: all B pairs in three launches, no host keypoint round trip
        m.ComputeStereoMatchesBatch(exL, exR, bf, fx)
        t0 = time.perf_counter()
        for _ in range(5):
            nmb, _, _ = m.ComputeStereoMatchesBatch(exL, exR, bf, fx)
        batch_s = (time.perf_counter() - t0) / 5 / B
        extract_pair_s = 1.0 / rateL + 1.0 / rateR
        pair_rate = 1.0 / (extract_pair_s + stereo_s)
        ba = alg_bytes(w, h, nf)
        cpu1 = cpu_rate([L[0], R[0]], nf, args.cpu_seconds / 2, 1) / 2
        cpuN = cpu_rate([p for pr in pairs for p in pr], nf, args.cpu_seconds, cores) / 2
        print(json.dumps({"config": "%dx%d stereo nf=%d" % (w, h, nf), "extract_pairs_per_s": 1.0 / extract_pair_s,
                          "stereo_match_ms_per_pair": stereo_s * 1e3, "pairs_per_s_with_stereo_match": pair_rate,
                          "stereo_match_batch_us_per_pair": batch_s * 1e6, "pairs_per_s_with_batched_stereo_match": 1.0 / (extract_pair_s + batch_s),
                          "stereo_matches_per_pair_batched": float(np.mean(nmb)),
                          "stereo_matches_per_pair": nm / nP, "keypoints_left": float(np.mean([len(k) for k, _ in resL])),
                          "alg_bytes_per_pair": 2 * ba, "hbm_frac_extract": 2 * ba / extract_pair_s / 1e9 / peak,
                          "cpu_pairs_per_s_1_thread": cpu1, "cpu_pairs_per_s_all_cores": cpuN, "cores": cores,
                          "note": "extraction device-resident, %d pairs per launch; stereo association called per pair through the C ABI "
                                  "(host keypoints in, synchronous)" % B}), flush=True)
        del exL, exR

    # ---- 1080p
    w, h, nf, B = 1920, 1080, 4000, 64
    frames = np.ascontiguousarray(np.stack([synth.rects(w, h, 5000 + (i % 4)) for i in range(B)]))
    rate, res, ex = gpu_extract_rate(frames, nf, steps=5)
    ok, od = O.OracleExtractor(nf, 1.2, 8, 20, 7)(frames[0])
    assert res[0][0].tobytes() == ok.tobytes() and np.array_equal(res[0][1], od)
    ba = alg_bytes(w, h, nf)
    cpu1 = cpu_rate([frames[0]], nf, args.cpu_seconds / 2, 1)
    cpuN = cpu_rate([frames[i] for i in range(4)], nf, args.cpu_seconds, cores)
    print(json.dumps({"config": "1920x1080 nf=4000", "frames_per_s": rate, "keypoints": float(np.mean([len(k) for k, _ in res])),
                      "alg_bytes_per_frame": ba, "hbm_frac": ba * rate / 1e9 / peak, "cpu_frames_per_s_1_thread": cpu1,
                      "cpu_frames_per_s_all_cores": cpuN, "cores": cores, "note": "device-resident, %d frames per launch" % B}), flush=True)


if __name__ == "__main__":
    main()
