#!/usr/bin/env python3
"""bench.py -- ORB front-end throughput on B200 (driver contract: see the task statement).

A step = one pass of the hot path (ORBextractor::operator(), reference src/ORBextractor.cc:1043-1105) over one batch of
B synthetic 640x480 frames with ORBextractor(1000, 1.2, 8, 20, 7) -- the configuration BASELINE.json's metric names.
  value  : frames/s, inputs resident in HBM when the timed region starts (orbx_extract_device), CUDA events on the
           handle's stream, max over ranks.
  e2e    : frames/s through the reference-facing C ABI with HOST buffers (orbx_extract_batch: pinned host images in,
           keypoints + descriptors out to host), host<->device copies inside the timed region.
  match  : brute-force Hamming Mpairs/s (10000x10000 and 2000x2000, SearchByBoW-shaped scan), device resident.
  roofline / cpu_baseline: see DESIGN.md section "Measurement".
`--impl reference` times the CPU restatement of the reference (oracle/, a scalar C++ port: the reference itself needs
OpenCV C++/ROS headers that this image does not have) on the host cores, same metric and config.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))

W, H, NF = 640, 480, 1000
PARAMS = (NF, 1.2, 8, 20, 7)
METRIC = "ORB extract frames/sec (640x480, 1000 kp)"
B_ALG = 5327340            # algorithmic bytes per 640x480 / 1000-kp frame (restated in DESIGN.md)
S_PYR = 950532             # sum of level pixels
S_PAD = 1158012            # sum of padded level pixels


def level_sizes():
    import numpy as np
    sf = [np.float32(1.0)]
    for _ in range(7):
        sf.append(np.float32(np.float64(sf[-1]) * np.float64(np.float32(1.2))))
    out = []
    for s in sf:
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(W) * inv)), int(np.rint(np.float32(H) * inv))))
    return out


def stage_alg_bytes(nkp):
    """Algorithmic bytes per frame of each stage (the split of B_ALG)."""
    return {
        "pyramid": W * H + S_PAD,                       # input read + padded pyramid write
        "fast_cells": S_PYR,                            # every level read once for detection
        "octree": 0,                                    # works on the candidate lists (no image bytes)
        "blur": 2 * S_PYR,                              # read + write of every level
        "orient_describe": min(S_PYR, 1369 * nkp) + 60 * nkp,   # blurred footprints + 28 B keypoint + 32 B descriptor
    }


def make_frames(n_distinct, seed0):
    import numpy as np
    from orbx import synth
    return np.stack([synth.rects(W, H, seed0 + i) for i in range(n_distinct)])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f:
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        self.f.close()
        os.unlink(self.f.name)
        if sm:
            sm.sort()
            out = {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}
        return out


def cpu_port_throughput(frames, seconds_budget, threads):
    """Oracle (scalar C++ port of the reference) on the host cores: one extractor instance and one frame per thread,
    as the reference runs (one thread per image).  Returns (frames/s, frames processed, threads)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py as O           # the one place bench.py executes oracle/: the measured CPU baseline
    O.lib()
    done = [0] * threads
    stop_at = [None]

    def worker(t):
        oe = O.OracleExtractor(*PARAMS)
        i = t
        while time.perf_counter() < stop_at[0]:
            oe(frames[i % len(frames)])
            done[t] += 1
            i += threads

    warm = O.OracleExtractor(*PARAMS)
    warm(frames[0])
    t0 = time.perf_counter()
    stop_at[0] = t0 + seconds_budget
    th = [threading.Thread(target=worker, args=(t,)) for t in range(threads)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    el = time.perf_counter() - t0
    return sum(done) / el, sum(done), threads


_REAL_STDOUT = None


def quiet_stdout():
    """The contract is ONE JSON line on stdout.  Libraries print there too (NCCL writes its version banner to stdout when the
    box sets NCCL_DEBUG): route file descriptor 1 to stderr for the run and keep the real stdout for the result line."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        os.write(1, data)
    else:
        os.write(_REAL_STDOUT, data)


PROFILE_TAG = "r01h"        # the committed ncu capture the traffic / instruction figures come from (profiles/<tag>_*)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    frames = make_frames(16, 1000)
    threads = os.cpu_count() or 1
    vals = []
    for _ in range(args.warmup):
        cpu_port_throughput(frames, 1.0, threads)
    per_step = max(2.0, min(10.0, 60.0 / max(args.steps, 1)))
    n = 0
    for _ in range(args.steps):
        v, cnt, _ = cpu_port_throughput(frames, per_step, threads)
        vals.append(v)
        n += cnt
    value = sum(vals) / len(vals)
    sample = "%d steps of %.1f s, %d frames in total, one oracle extractor + one frame per thread" % (args.steps, per_step, n)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": "ORBextractor(1000,1.2,8,20,7) on synthetic 640x480 'rects' frames (BASELINE.json configs[0])"},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def pin_to_gpu_numa_node(local):
    """Bind this rank (and the host threads it starts) to the CPUs next to its GPU, so that the pinned staging buffers are
    allocated on the GPU's own NUMA node: with 8 ranks per box the host side of the e2e path is otherwise limited by
    cross-socket traffic.  Returns a description for the JSON line."""
    try:
        import torch
        p = torch.cuda.get_device_properties(local)
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        txt = open("/sys/bus/pci/devices/%s/local_cpulist" % bdf).read().strip()
        cpus = set()
        for part in txt.split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return "%s cpus %s" % (bdf, txt)
    except Exception as e:      # not fatal: the run proceeds unpinned
        return "unpinned (%s)" % type(e).__name__
    return "unpinned"


def run_ours(args):
    import numpy as np
    import torch
    import orbx

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    numa = pin_to_gpu_numa_node(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    B = args.batch
    distinct = make_frames(32, 1000 + 1000 * rank)                 # frames are independent: every rank has its own
    host = np.ascontiguousarray(np.tile(distinct, (B // 32 + 1, 1, 1))[:B])
    dev = torch.from_numpy(host).cuda()                            # B x 480 x 640 u8, resident in HBM (larger than L2 for B >= 512)
    ex = orbx.ORBextractor(*PARAMS, device=local, max_batch=B)
    stream = torch.cuda.ExternalStream(ex.stream())
    torch.cuda.synchronize()

    def step():
        ex.extract_device(dev.data_ptr(), B, W, H, W, W * H)

    # parity gate before any number is taken: the first frames must equal the CPU oracle byte for byte
    step()
    res = ex.fetch()
    parity = "unchecked"
    nkp_mean = float(np.mean([len(k) for k, _ in res]))
    if rank == 0:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import oracle_py as O       # checker only
        oe = O.OracleExtractor(*PARAMS)
        for b in range(2):
            ok, od = oe(host[b])
            assert res[b][0].tobytes() == ok.tobytes() and np.array_equal(res[b][1], od), "GPU result differs from the oracle"
        parity = "bit-exact vs oracle on 2 frames"

    for _ in range(max(args.warmup, 3)):
        step()
    ex.sync()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    e1.synchronize()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    launches = ex.last_launches() * args.steps
    # the timed region lasts tens of milliseconds, shorter than nvidia-smi's sampling period: the identical steps keep running
    # (untimed) for one more second so that the sampler sees the clocks this workload runs at
    t_soak = time.perf_counter()
    while time.perf_counter() - t_soak < 1.0:
        step()
        ex.sync()
    clk = clocks.stop()
    clk["window"] = "timed steps + 1 s of the same steps, untimed"
    # per-stage device time (events between the launches, same stream), a few extra steps outside the timed region
    ex.set_profiling(True)
    acc = {}
    for _ in range(3):
        step()
        for k, v in ex.stage_ms().items():
            acc[k] = acc.get(k, 0.0) + v / 3
    ex.set_profiling(False)
    if dist is not None:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = B * world / (ms_per_step / 1e3)

    # ---- e2e: host buffers through orbx_extract_batch, one handle per host thread so that H2D, kernels and D2H overlap
    # (measured sweep, frames/s: 4 x 128 frames 164 k, 4 x 64 170 k, 6 x 64 171 k, 8 x 64 171 k, 2 x 256 159 k)
    import ctypes as C
    L = orbx.lib()
    if args.e2e_wc:      # frames in write-combined page-locked memory: the host only writes them, the copy engine reads them unsnooped
        pinned_in = orbx.PinnedArray(host.shape, np.uint8, write_combined=True)
        pinned_in.array[...] = host
        base_ptr = pinned_in.ptr
    else:
        pinned = torch.from_numpy(host).pin_memory()
        base_ptr = pinned.data_ptr()
    chunk = min(args.e2e_chunk, B)
    nthreads = args.e2e_threads
    workers = [orbx.ORBextractor(*PARAMS, device=local, max_batch=chunk) for _ in range(nthreads)]
    for wk in workers:
        wk(host[0])                                                  # plans the shape; max_keypoints is exact afterwards
    cap = workers[0].max_keypoints()
    # result arrays of the caller: page-locked and laid out with cap = the handle's per-frame capacity, so that the library
    # copies keypoints and descriptors straight into them (no staging, no host-side unpacking)
    pin_k = [torch.zeros((chunk, cap, 28), dtype=torch.uint8).pin_memory() for _ in range(nthreads)]
    pin_d = [torch.zeros((chunk, cap, 32), dtype=torch.uint8).pin_memory() for _ in range(nthreads)]
    outs = [(pin_k[t].numpy().view(orbx.KP_DTYPE).reshape(chunk, cap), pin_d[t].numpy(), np.zeros(chunk, np.int32)) for t in range(nthreads)]
    chunks = [(c0, min(chunk, B - c0)) for c0 in range(0, B, chunk)]

    def e2e_worker(t, nsteps):
        torch.cuda.set_device(local)
        k, d, n = outs[t]
        for _ in range(nsteps):
            for ci in range(t, len(chunks), nthreads):
                c0, cb = chunks[ci]
                ptrs = (C.c_void_p * cb)(*[base_ptr + (c0 + i) * W * H for i in range(cb)])
                rc = L.orbx_extract_batch(workers[t]._h, ptrs, cb, W, H, W, k.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p), cap,
                                          n.ctypes.data_as(C.c_void_p))
                if rc != 0:
                    raise RuntimeError(L.orbx_last_error().decode())

    def e2e_run(nsteps):
        th = [threading.Thread(target=e2e_worker, args=(t, nsteps)) for t in range(nthreads)]
        for x in th:
            x.start()
        for x in th:
            x.join()

    e2e_run(1)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e2e_run(args.steps)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = B * world * args.steps / e2e_s
    h2d = B * W * H
    d2h = B * (cap * 60 + 4)

    # ---- matcher: brute-force Hamming, device resident
    match = {}
    if rank == 0:
        from orbx import synth
        m = orbx.ORBmatcher(0.9, True, device=local)
        ms_ = torch.cuda.ExternalStream(m.stream())
        for N in (2000, 10000):
            A, Bd = synth.descriptor_sets(N, 4000 + N)
            dq, dt = torch.from_numpy(A).cuda(), torch.from_numpy(Bd).cuda()
            o = [torch.zeros(N, dtype=torch.int32, device="cuda") for _ in range(3)]
            torch.cuda.synchronize()
            reps = 20
            for _ in range(3):
                m.match_bruteforce_device(dq.data_ptr(), N, dt.data_ptr(), N, *[x.data_ptr() for x in o])
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(ms_)
            for _ in range(reps):
                m.match_bruteforce_device(dq.data_ptr(), N, dt.data_ptr(), N, *[x.data_ptr() for x in o])
            a1.record(ms_)
            a1.synchronize()
            t_ms = a0.elapsed_time(a1) / reps
            match["%dx%d" % (N, N)] = {"Mpairs_per_s": N * N / (t_ms / 1e3) / 1e6, "ms": t_ms,
                                        "popc32_per_s": 8.0 * N * N / (t_ms / 1e3)}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    alg = stage_alg_bytes(NF)
    top = max(acc, key=acc.get)
    top_bytes = alg[top] if alg[top] > 0 else alg["fast_cells"]
    achieved = top_bytes * B / (acc[top] / 1e3) / 1e9
    pipeline_gbs = B_ALG * value / world / 1e9
    traffic, traffic_src, issue = None, None, None
    try:      # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture (same shape), scaled to B frames
        tr = json.load(open(os.path.join(ROOT, "profiles", PROFILE_TAG + "_traffic.json")))
        per_frame = sum(tr["dram_bytes_per_frame"][k] for k in top.split("+"))
        traffic, traffic_src = per_frame * B, tr["source"]
        # the resource that does bind this integer path: warp-instruction issue slots (148 SMs x 4 schedulers x SM clock);
        # instruction counts per frame from the same capture, rate = this run's frames/s
        inst = tr.get("warp_instructions_per_frame")
        if inst:
            smclk = float(peaks.get("sm_max_mhz", 1965.0)) * 1e6
            issue_peak = 148 * 4 * smclk
            issue = {"warp_instructions_per_frame": inst, "total_per_frame": sum(inst.values()),
                     "achieved_warp_inst_per_s": sum(inst.values()) * value / world, "peak_warp_inst_per_s": issue_peak,
                     "frac": sum(inst.values()) * value / world / issue_peak,
                     "note": "issue-slot utilisation of the whole extraction (ncu instruction counts x measured frames/s); "
                             "north_star names the HBM roofline, ncu shows the integer pipes bind"}
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": top_bytes * B,
                "stage_ms": acc, "stage_alg_bytes_per_frame": alg,
                "pipeline": {"alg_bytes_per_frame": B_ALG, "achieved": pipeline_gbs, "frac": pipeline_gbs / peak,
                             "note": "whole extraction per GPU: B_alg x frames/s / peak"},
                "issue_slots": issue}
    cpu = None
    if world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        v, cnt, _ = cpu_port_throughput(host[:16], 12.0, threads)
        v1, cnt1, _ = cpu_port_throughput(host[:16], 4.0, 1)
        cpu = {"value": v, "unit": "frames/s", "cores": threads, "kind": "port",
               "sample": "%d frames in 12 s, one oracle extractor + one frame per thread (scalar C++ port of the reference; "
                         "the reference itself needs OpenCV C++/ROS headers absent here)" % cnt,
               "single_thread_value": v1}
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "ORBextractor(1000,1.2,8,20,7) on synthetic 640x480 'rects' frames (BASELINE.json configs[0], the "
                               "configuration the metric names)", "frames_per_step_per_gpu": B, "parallelism": "frames sharded per GPU, no collective",
                   "l2": "inputs larger than L2 (%d MB per step)" % (B * W * H >> 20), "keypoints_per_frame": nkp_mean},
        "clocks": clk, "gpu_launches": launches, "parity": parity,
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "how": "orbx_extract_batch on pinned host frames into pinned host results, %d handles x %d-frame chunks on %d host threads" % (nthreads, chunk, nthreads),
                "numa": numa},
        "roofline": roofline, "cpu_baseline": cpu, "match": match,
    }
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=512, help="frames per step per GPU")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--e2e-threads", type=int, default=6, help="host threads (= extractor handles) of the e2e leg")
    ap.add_argument("--e2e-wc", type=int, default=0, help="1: e2e input frames in write-combined page-locked memory")
    ap.add_argument("--e2e-chunk", type=int, default=64, help="frames per orbx_extract_batch call in the e2e leg")
    args = ap.parse_args()
    quiet_stdout()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
