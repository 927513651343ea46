#!/usr/bin/env python3
"""bench.py -- ORB front-end throughput on B200 (driver contract: see the task statement).

A step = one pass of the hot path (ORBextractor::operator(), reference src/ORBextractor.cc:1043-1105; for the stereo configurations
also Frame::ComputeStereoMatches, src/Frame.cc:481-655) over one batch of synthetic input.  `--config` picks the BASELINE.json
configuration: 0 (default, the one the metric names) = 640x480 / 1000 features, 1 = 752x480 stereo / 1200, 2 = 1241x376 stereo /
2000, 4 = 1920x1080 / 4000 (config 3 is the matcher alone: the `match` object of every line).
  value  : frames/s (stereo: pairs/s), inputs resident in HBM when the timed region starts (orbx_extract_device), CUDA events,
           max over ranks.
  e2e    : the same through the reference-facing C ABI with HOST buffers (orbx_extract_batch: pinned host images in, keypoints +
           descriptors [+ uRight / depth] out to host), host<->device copies inside the timed region.
  match  : brute-force Hamming Mpairs/s (10000x10000 and 2000x2000, SearchByBoW-shaped scan), device resident.
  roofline / cpu_baseline: see DESIGN.md section "Measurement".
`--impl reference` times the reference's own CPU code on the host cores: oracle/_ref, the unmodified src/ORBextractor.cc compiled
over a cv shim (kind "reference"); the stereo configurations run the reference's own Frame per pair (oracle/_ref/libframe_ref.so: the
unmodified src/Frame.cc with ExtractORB on both images and ComputeStereoMatches).  Same metric, config and workload string as the CUDA arm.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "orb_slam2v2-1_b200"))

# BASELINE.json configs: shape, nfeatures, stereo calibration (upstream EuRoC / KITTI values), frames or
# pairs per step per GPU (inputs larger than L2 in every case), seed of the first synthetic frame
CONFIGS = {
    0: dict(W=640, H=480, NF=1000, stereo=False, B=512, seed=1000, name="BASELINE.json configs[0]"),
    1: dict(W=752, H=480, NF=1200, stereo=True, B=192, seed=2000, fx=435.2, bf=47.9, name="BASELINE.json configs[1]"),
    2: dict(W=1241, H=376, NF=2000, stereo=True, B=128, seed=3000, fx=718.856, bf=386.1448, name="BASELINE.json configs[2]"),
    4: dict(W=1920, H=1080, NF=4000, stereo=False, B=64, seed=5000, name="BASELINE.json configs[4]"),
}
PROFILE_TAG = "r02"         # profiles/<tag>_traffic.json: ncu counters, stamped with the kernel sources they were taken from


def params(cfg):
    return (cfg["NF"], 1.2, 8, 20, 7)


def metric_of(cfg):
    if cfg["stereo"]:
        return "ORB extract + stereo match pairs/sec (%dx%d stereo, %d kp)" % (cfg["W"], cfg["H"], cfg["NF"])
    return "ORB extract frames/sec (%dx%d, %d kp)" % (cfg["W"], cfg["H"], cfg["NF"])


def unit_of(cfg):
    return "pairs/s" if cfg["stereo"] else "frames/s"


def workload_of(cfg):
    """One string for both arms (the driver compares them)."""
    what = "stereo pairs: both images extracted, then Frame::ComputeStereoMatches" if cfg["stereo"] else "'rects' frames"
    return "ORBextractor(%d,1.2,8,20,7) on synthetic %dx%d %s (%s)" % (cfg["NF"], cfg["W"], cfg["H"], what, cfg["name"])


def level_sizes(W, H):
    """Pyramid geometry: float chain mvScaleFactor[i] = mvScaleFactor[i-1] * 1.2f, size = cvRound((float)dim * inv) (ORBextractor.cc:421,1112)."""
    import numpy as np
    sf = [np.float32(1.0)]
    for _ in range(7):
        sf.append(np.float32(np.float64(sf[-1]) * np.float64(np.float32(1.2))))
    out = []
    for s in sf:
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(W) * inv)), int(np.rint(np.float32(H) * inv))))
    return out


def stage_alg_bytes(cfg):
    """Algorithmic bytes per IMAGE of each stage; their sum is B_alg of SURVEY.md section 8(d):
    W*H + Spad + S + 2S + min(S, 1369*N) + 60*N (640x480 / 1000: 5 327 340)."""
    W, H, N = cfg["W"], cfg["H"], cfg["NF"]
    lv = level_sizes(W, H)
    S = sum(w * h for w, h in lv)
    Spad = sum((w + 38) * (h + 38) for w, h in lv)
    return {
        "pyramid": W * H + Spad,                        # input read + padded pyramid write
        "fast_cells": S,                                # every level read once for detection
        "octree": 0,                                    # works on the candidate lists (no image bytes)
        "blur": 2 * S,                                  # read + write of every level
        "orient_describe": min(S, 1369 * N) + 60 * N,   # blurred footprints + 28 B keypoint + 32 B descriptor
    }


def alg_bytes_per_unit(cfg):
    b = sum(stage_alg_bytes(cfg).values())
    return 2 * b if cfg["stereo"] else b


def make_inputs(cfg, n_distinct, seed0):
    """[n, H, W] left (or only) frames and, for stereo, the right frames."""
    import numpy as np
    from orbx import synth
    W, H = cfg["W"], cfg["H"]
    if cfg["stereo"]:
        pairs = [synth.stereo_rects(W, H, seed0 + i)[:2] for i in range(n_distinct)]
        return np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])
    return np.stack([synth.rects(W, H, seed0 + i) for i in range(n_distinct)]), None


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


def kernel_source_sha16():
    """Stamp that ties profiles/<tag>_traffic.json (ncu counters) to the kernel sources they were measured on."""
    h = hashlib.sha256()
    for f in ("extract_kernels.cu", "extract_kernels.cuh", "orbx_extractor.cu"):
        h.update(open(os.path.join(ROOT, "orb_slam2v2-1_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


# ------------------------------------------------------------------------------------------------ CPU arm (checker code, timed)
def _checker():
    """(module, kind): oracle/_ref -- the unmodified reference extractor -- when its libraries exist, else the oracle port."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import ref_py as R          # this and cpu_reference_throughput are the places bench.py executes oracle/: the measured CPU baseline
    if R.available():
        return R, "reference"
    import oracle_py as O
    return O, "port"


def _cpu_extractor(cfg, allocator):
    mod, kind = _checker()
    if kind == "reference":
        return mod.RefExtractor(*params(cfg), allocator=allocator, variant="fma")      # the reference's own flags
    return mod.OracleExtractor(*params(cfg), tie_mode=0, fma_mode=1)


def _stereo_pipeline(cfg):
    """The reference's own Frame for one rectified pair (oracle/_ref/libframe_ref.so: unmodified Frame.cc + ORBextractor.cc + ORBmatcher.cc):
    extraction of both images, grid, ComputeStereoMatches; None where that library is not built (then cpu_unit's port is used)."""
    _, kind = _checker()
    if kind != "reference":
        return None
    import ref_frame_py as RF
    if not RF.available():
        return None
    return RF.RefStereoPipeline([cfg["fx"], cfg["fx"], cfg["W"] * 0.5, cfg["H"] * 0.5], cfg["bf"], 35.0, *params(cfg))


def cpu_unit(cfg, exL, exR, left, right):
    """One unit of the workload on the CPU, as the reference runs it: operator() on the image (stereo: on both, then
    Frame::ComputeStereoMatches on the two mvImagePyramids -- the oracle's port of it, which tests/test_reference_frame.py holds equal to
    the reference's; the timed legs use the reference's own Frame instead when libframe_ref.so is there, see _stereo_pipeline)."""
    kl, dl = exL(left)
    if not cfg["stereo"]:
        return kl, dl
    import oracle_py as O
    kr, dr = exR(right)
    pyrL = [exL.level_padded(l)[19:-19, 19:-19] for l in range(8)]
    pyrR = [exR.level_padded(l)[19:-19, 19:-19] for l in range(8)]
    sf, isf = exL.scale_tables()[:2]
    n, uR, dep = O.stereo_matches(kl, dl, kr, dr, pyrL, pyrR, sf, isf, cfg["bf"], cfg["fx"])
    return kl, dl, kr, dr, n, uR, dep


def cpu_reference_throughput(cfg, left, right, seconds_budget, threads):
    """Reference CPU code on the host cores: one extractor instance (two for stereo) and one unit per thread, plain malloc as a
    stock build.  ctypes releases the GIL inside the calls.  Returns (units/s, units processed)."""
    done = [0] * threads
    errors = []
    stop_at = [None]

    def worker(t):
        try:
            pipe = _stereo_pipeline(cfg) if cfg["stereo"] else None
            if pipe is None:
                exL = _cpu_extractor(cfg, 0)
                exR = _cpu_extractor(cfg, 0) if cfg["stereo"] else None
            i = t
            while time.perf_counter() < stop_at[0]:
                if pipe is not None:
                    pipe.run(left[i % len(left)], right[i % len(left)])
                else:
                    cpu_unit(cfg, exL, exR, left[i % len(left)], right[i % len(left)] if cfg["stereo"] else None)
                done[t] += 1
                i += threads
        except Exception as e:          # noqa: a dead worker must not turn into a silently higher or lower number
            errors.append(e)

    warmL = _cpu_extractor(cfg, 0)
    cpu_unit(cfg, warmL, _cpu_extractor(cfg, 0) if cfg["stereo"] else None, left[0], right[0] if cfg["stereo"] else None)
    t0 = time.perf_counter()
    stop_at[0] = t0 + seconds_budget
    th = [threading.Thread(target=worker, args=(t,)) for t in range(threads)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    if errors:
        raise errors[0]
    el = time.perf_counter() - t0
    return sum(done) / el, sum(done)


def cv2_primitives_ms(cfg, img):
    """How far is the scalar cv shim under oracle/_ref from SIMD OpenCV?  The three OpenCV primitives that dominate the reference's
    extractor, on one frame with the real cv2 wheel, one thread: 7 resizes, cell-wise FAST over every level, 8 blurs.  None when cv2
    is not importable."""
    try:
        import cv2
    except Exception:
        return None
    cv2.setNumThreads(1)
    lv = level_sizes(cfg["W"], cfg["H"])
    fast_ini = cv2.FastFeatureDetector_create(20, True)

    def once():
        t = {}
        t0 = time.perf_counter()
        pyr = [img]
        for (w, h) in lv[1:]:
            pyr.append(cv2.resize(pyr[-1], (w, h), interpolation=cv2.INTER_LINEAR))
        t["resize"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        for p in pyr:
            hh, ww = p.shape
            for y in range(16, hh - 19, 30):
                for x in range(16, ww - 19, 30):
                    fast_ini.detect(p[y:min(y + 36, hh - 16), x:min(x + 36, ww - 16)])
        t["fast_cells"] = time.perf_counter() - t0
        t0 = time.perf_counter()
        for p in pyr:
            cv2.GaussianBlur(p, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)
        t["blur"] = time.perf_counter() - t0
        return t
    once()
    runs = [once() for _ in range(3)]
    out = {k: min(r[k] for r in runs) * 1e3 for k in runs[0]}
    out["sum"] = sum(out.values())
    return out


def cpu_baseline_object(cfg, left, right, seconds, with_single=True):
    threads = os.cpu_count() or 1
    _, kind = _checker()
    v, cnt = cpu_reference_throughput(cfg, left, right, seconds, threads)
    what = ("unmodified reference src/ORBextractor.cc (oracle/_ref, g++ -O3 -march=x86-64-v3, cv primitives = the cv2-pinned scalar shim)"
            if kind == "reference" else "scalar C++ port of the reference (oracle/)")
    if cfg["stereo"]:
        what += (" inside the unmodified reference src/Frame.cc (constructor + ExtractORB(1, right) + ComputeStereoMatches, libframe_ref.so)"
                 if _stereo_pipeline(cfg) is not None else " + Frame::ComputeStereoMatches from the oracle port")
    out = {"value": v, "unit": unit_of(cfg), "cores": threads, "kind": kind,
           "sample": "%d %s in %.0f s, one extractor instance%s and one unit per thread; %s" %
                     (cnt, "pairs" if cfg["stereo"] else "frames", seconds, " pair" if cfg["stereo"] else "", what)}
    if with_single:
        v1, _ = cpu_reference_throughput(cfg, left, right, max(seconds / 3, 2.0), 1)
        out["single_thread_value"] = v1
        out["cv2_primitives_ms"] = cv2_primitives_ms(cfg, left[0])
    return out


def run_reference(args, cfg):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    left, right = make_inputs(cfg, 8 if cfg["W"] > 1500 else 16, cfg["seed"])
    threads = os.cpu_count() or 1
    _, kind = _checker()
    for _ in range(args.warmup):
        cpu_reference_throughput(cfg, left, right, 1.0, threads)
    per_step = max(0.5, min(10.0, 60.0 / max(args.steps, 1)))          # about a minute of CPU work whatever K is
    vals, n = [], 0
    for _ in range(args.steps):
        v, cnt = cpu_reference_throughput(cfg, left, right, per_step, threads)
        vals.append(v)
        n += cnt
    value = sum(vals) / len(vals)
    sample = "%d steps of %.1f s, %d units in total, one reference extractor%s + one unit per thread" % (
        args.steps, per_step, n, " pair" if cfg["stereo"] else "")
    if cfg["stereo"] and _stereo_pipeline(cfg) is not None:
        sample += "; each pair through the unmodified reference src/Frame.cc (constructor + ExtractORB(1, right) + ComputeStereoMatches)"
    emit({
        "impl": "reference", "metric": metric_of(cfg), "value": value, "unit": unit_of(cfg), "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic", "config": {"workload": workload_of(cfg)},
        "cpu_baseline": {"value": value, "unit": unit_of(cfg), "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": unit_of(cfg), "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


# ------------------------------------------------------------------------------------------------ CUDA arm
def pin_to_gpu_numa_node(local):
    """Bind this rank (and the host threads it starts) to the CPUs next to its GPU, so that the pinned staging buffers are
    allocated on the GPU's own NUMA node.  Returns a description for the JSON line."""
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


def parity_gate(cfg, hostL, hostR, resL, resR, stereo_out, frames):
    frames = min(frames, len(resL))
    """Before any number is taken: this rank's first `frames` units must equal the reference build byte for byte (keypoints,
    descriptors; stereo: uRight / depth of the first two pairs).  The CUDA path runs with fma_mode 1 = the reference's own flags."""
    import numpy as np
    _, kind = _checker()
    exL = _cpu_extractor(cfg, 1)            # monotone node allocator: the definition the quadtree kernel is held to
    exR = _cpu_extractor(cfg, 1) if cfg["stereo"] else None
    B = len(resL)
    # half of the checked units from the start of the batch, half from its end (a large batch runs as two halves on two streams)
    picks = list(range((frames + 1) // 2)) + list(range(B - frames // 2, B))
    for b in picks:
        if cfg["stereo"] and b < 2:
            kl, dl, kr, dr, n, uR, dep = cpu_unit(cfg, exL, exR, hostL[b], hostR[b])
            nm, guR, gdep = stereo_out
            assert nm[b] == n and guR[b, :len(kl)].tobytes() == uR.tobytes() and gdep[b, :len(kl)].tobytes() == dep.tobytes(), \
                "stereo association of pair %d differs from the oracle" % b
        else:
            kl, dl = exL(hostL[b])
            kr, dr = exR(hostR[b]) if cfg["stereo"] else (None, None)
        assert resL[b][0].tobytes() == kl.tobytes() and np.array_equal(resL[b][1], dl), "GPU result of frame %d differs from the %s" % (b, kind)
        if cfg["stereo"]:
            assert resR[b][0].tobytes() == kr.tobytes() and np.array_equal(resR[b][1], dr), "GPU result of right frame %d differs from the %s" % (b, kind)
    return "bit-exact vs %s on %d %s of this rank" % ("oracle/_ref (the unmodified reference extractor)" if kind == "reference" else "the oracle port",
                                                     frames, "pairs" if cfg["stereo"] else "frames")


def padded_device(host, pitch):
    import torch
    B, H, W = host.shape
    if pitch == W:
        return torch.from_numpy(host).cuda()
    dev = torch.zeros((B, H, pitch), dtype=torch.uint8, device="cuda")
    dev[:, :, :W] = torch.from_numpy(host).cuda()
    return dev


def run_ours(args, cfg):
    import ctypes as C
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

    W, H, stereo = cfg["W"], cfg["H"], cfg["stereo"]
    B = args.batch or cfg["B"]
    nd = 8 if W > 1500 else (16 if stereo else 32)
    dL, dR = make_inputs(cfg, nd, cfg["seed"] + 1000 * rank)       # units are independent: every rank has its own
    hostL = np.ascontiguousarray(np.tile(dL, (B // nd + 1, 1, 1))[:B])
    hostR = np.ascontiguousarray(np.tile(dR, (B // nd + 1, 1, 1))[:B]) if stereo else None
    pitch = (W + 15) & ~15                                         # rows on 16-byte boundaries: what a caller who keeps frames in HBM would do (the TMA
                                                                   # staging needs it; 4-byte aligned rows are accepted and take the LDGSTS kernels)
    devL = padded_device(hostL, pitch)                             # B x H x pitch u8, resident in HBM, larger than L2
    devR = padded_device(hostR, pitch) if stereo else None
    exL = orbx.ORBextractor(*params(cfg), device=local, max_batch=B)
    exR = orbx.ORBextractor(*params(cfg), device=local, max_batch=B) if stereo else None
    for e in (exL, exR):
        if e is not None:
            e.set_fma_mode(1)                                      # ORBextractor.cc:119-120 as the reference's own flags compile it
    m = orbx.ORBmatcher(0.6, True, device=local)
    L = orbx.lib()
    stream0 = torch.cuda.ExternalStream(exL.stream())
    stream1 = torch.cuda.ExternalStream(m.stream()) if stereo else stream0
    torch.cuda.synchronize()
    scap = [None]
    sout = [None]

    def step():
        exL.extract_device(devL.data_ptr(), B, W, H, pitch, pitch * H)
        if stereo:
            exR.extract_device(devR.data_ptr(), B, W, H, pitch, pitch * H)
            if scap[0] is None:
                exL.sync()
                cap = max(exL.max_keypoints(), 64)
                scap[0] = (cap, np.zeros((B, cap), np.float32), np.zeros((B, cap), np.float32), np.zeros(B, np.int32))
            cap, uR, dep, nm = scap[0]
            rc = L.orbm_stereo_matches_batch(m._h, exL._h, exR._h, cfg["bf"], 0.0, cfg["fx"], uR.ctypes.data_as(C.c_void_p),
                                             dep.ctypes.data_as(C.c_void_p), cap, nm.ctypes.data_as(C.c_void_p))
            if rc != 0:
                raise RuntimeError(L.orbx_last_error().decode())
            sout[0] = (nm, uR, dep)

    step()
    resL = exL.fetch()
    resR = exR.fetch() if stereo else None
    nkp_mean = float(np.mean([len(k) for k, _ in resL]))
    parity = parity_gate(cfg, hostL, hostR, resL, resR, sout[0], min(args.parity_frames, B))

    for _ in range(max(args.warmup, 3)):
        step()
    exL.sync()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = ClockSampler(local)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream0)
    for _ in range(args.steps):
        step()
    e1.record(stream1)          # stereo: the association call waits for both extractions, so its stream sees the end of the step
    e1.synchronize()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    per_step_launches = exL.last_launches() * (2 if stereo else 1) + (L.orbm_last_launches(m._h) if stereo else 0)
    launches = per_step_launches * args.steps
    # the timed region lasts tens of milliseconds, shorter than nvidia-smi's sampling period: the identical steps keep running
    # (untimed) for one more second so that the sampler sees the clocks this workload runs at
    t_soak = time.perf_counter()
    while time.perf_counter() - t_soak < 1.0:
        step()
        exL.sync()
    clk = clocks.stop()
    clk["window"] = "timed steps + 1 s of the same steps, untimed"
    # per-stage device time (events between the launches, same stream), a few extra steps outside the timed region
    exL.set_profiling(True)
    acc = {}
    for _ in range(3):
        exL.extract_device(devL.data_ptr(), B, W, H, pitch, pitch * H)
        for k, v in exL.stage_ms().items():
            acc[k] = acc.get(k, 0.0) + v / 3
    exL.set_profiling(False)
    if dist is not None:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = B * world / (ms_per_step / 1e3)
    return finish_ours(args, cfg, dict(locals()))


def e2e_leg(args, cfg, S):
    """Host buffers through the C ABI: pinned host frames in (laid out like the device slab, so a chunk travels as one linear copy),
    pinned host keypoints + descriptors out, stereo uRight / depth out; one handle set per host thread so that H2D, kernels and D2H
    of different chunks overlap.  Returns (units/s, h2d bytes per step, d2h bytes per step, description)."""
    import ctypes as C
    import numpy as np
    import torch
    import orbx
    L, B, W, H, stereo, local, dist, world = S["L"], S["B"], cfg["W"], cfg["H"], cfg["stereo"], S["local"], S["dist"], S["world"]
    P32 = (W + 31) & ~31
    frame = (P32 * H + 255) & ~255
    sides = [S["hostL"]] + ([S["hostR"]] if stereo else [])
    pinned = []
    for host in sides:
        if args.e2e_wc:      # frames in write-combined page-locked memory: the host only writes them, the copy engine reads them unsnooped
            pa = orbx.PinnedArray((B, frame), np.uint8, write_combined=True)
            arr, base = pa.array, pa.ptr
        else:
            pa = torch.zeros((B, frame), dtype=torch.uint8).pin_memory()
            arr, base = pa.numpy(), pa.data_ptr()
        arr[:, :P32 * H].reshape(B, H, P32)[:, :, :W] = host
        pinned.append((pa, base))
    chunk = min(args.e2e_chunk, B)
    nthreads = args.e2e_threads
    # what the box can move: the same pinned frames copied to the device and nothing else, all ranks at once -> the e2e ceiling
    dst = torch.empty((B, frame), dtype=torch.uint8, device="cuda")
    src_t = [torch.from_numpy(np.frombuffer((C.c_uint8 * (B * frame)).from_address(base), dtype=np.uint8).reshape(B, frame)) for _, base in pinned]
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    c0e, c1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dst.copy_(src_t[0], non_blocking=True)
    c0e.record()
    for _ in range(8):
        for t_ in src_t:
            dst.copy_(t_, non_blocking=True)
    c1e.record()
    c1e.synchronize()
    h2d_gbs = 8 * len(sides) * B * frame / (c0e.elapsed_time(c1e) / 1e3) / 1e9
    del dst
    errors = []
    if not stereo:
        # the library's own driver (orbx_pool_*): this rank's GPU, `nthreads` lanes (host threads with their own handles) that take
        # chunks of the batch in turn; results go straight into the caller's page-locked arrays at the frame's index
        pool = orbx.ExtractorPool(*params(cfg), devices=[local], chunk=chunk, lanes=nthreads)
        pool.set_fma_mode(1)
        cap = pool.plan(W, H)
        scapacity = cap
        # two steps in flight (orbx_pool_submit_batch / orbx_pool_wait), each with its own result arrays: the lanes go from the last
        # chunks of one step to the first of the next without the fill / drain gap of a synchronous call per step -- what the stereo
        # leg's worker threads do by themselves.  Every step's H2D, kernels and D2H lie inside the timed region.
        DEPTH = 2
        outs_ = [(torch.zeros((B, cap, 28), dtype=torch.uint8).pin_memory(), torch.zeros((B, cap, 32), dtype=torch.uint8).pin_memory(),
                  np.zeros(B, np.int32)) for _ in range(DEPTH)]
        pk, pd, nn = outs_[0]
        ptrs = (C.c_void_p * B)(*[pinned[0][1] + i * frame for i in range(B)])

        def run(nsteps):
            tickets = []
            for s_ in range(nsteps):
                if len(tickets) == DEPTH:
                    if L.orbx_pool_wait(pool._h, tickets.pop(0)) != 0:
                        raise RuntimeError(L.orbx_last_error().decode())
                k_, d_, n_ = outs_[s_ % DEPTH]
                t_ = C.c_ulonglong(0)
                rc = L.orbx_pool_submit_batch(pool._h, ptrs, B, W, H, P32, C.c_void_p(k_.data_ptr()), C.c_void_p(d_.data_ptr()), cap,
                                              n_.ctypes.data_as(C.c_void_p), C.byref(t_))
                if rc != 0:
                    raise RuntimeError(L.orbx_last_error().decode())
                tickets.append(t_.value)
            for t_ in tickets:
                if L.orbx_pool_wait(pool._h, t_) != 0:
                    raise RuntimeError(L.orbx_last_error().decode())

        run(1)
        # the pool's results are the device-resident leg's: same frames, same bytes
        resL = S["resL"]
        for b in (0, B // 2, B - 1):
            k = pk[b].numpy().view(orbx.KP_DTYPE).reshape(cap)[:nn[b]]
            assert k.tobytes() == resL[b][0].tobytes() and np.array_equal(pd[b].numpy()[:nn[b]], resL[b][1]), "pool result differs (frame %d)" % b
    else:
        workers = []
        for _ in range(nthreads):
            hs = [orbx.ORBextractor(*params(cfg), device=local, max_batch=chunk) for _ in sides]
            for h, host in zip(hs, sides):
                h.set_fma_mode(1)
                h(host[0])                                                  # plans the shape; max_keypoints is exact afterwards
            workers.append((hs, orbx.ORBmatcher(0.6, True, device=local)))
        cap = workers[0][0][0].max_keypoints()
        scapacity = max(cap, 64)
        outs = []
        for _ in range(nthreads):
            per_side = []
            for _ in sides:   # result arrays of the caller: page-locked and laid out with cap = the handle's per-frame capacity, so
                pk = torch.zeros((chunk, cap, 28), dtype=torch.uint8).pin_memory()      # that the library copies straight into them
                pd = torch.zeros((chunk, cap, 32), dtype=torch.uint8).pin_memory()
                per_side.append((pk, pd, np.zeros(chunk, np.int32)))
            outs.append((per_side, np.zeros((chunk, scapacity), np.float32), np.zeros((chunk, scapacity), np.float32), np.zeros(chunk, np.int32)))
        chunks = [(c0, min(chunk, B - c0)) for c0 in range(0, B, chunk)]

        def worker(t, nsteps):
            try:
                torch.cuda.set_device(local)
                hs, m = workers[t]
                per_side, uR, dep, nm = outs[t]
                for _ in range(nsteps):
                    for ci in range(t, len(chunks), nthreads):
                        c0, cb = chunks[ci]
                        for h, (_, base), (pk, pd, n) in zip(hs, pinned, per_side):
                            ptrs = (C.c_void_p * cb)(*[base + (c0 + i) * frame for i in range(cb)])
                            rc = L.orbx_extract_batch(h._h, ptrs, cb, W, H, P32, C.c_void_p(pk.data_ptr()), C.c_void_p(pd.data_ptr()), cap,
                                                      n.ctypes.data_as(C.c_void_p))
                            if rc != 0:
                                raise RuntimeError(L.orbx_last_error().decode())
                        rc = L.orbm_stereo_matches_batch(m._h, hs[0]._h, hs[1]._h, cfg["bf"], 0.0, cfg["fx"], uR.ctypes.data_as(C.c_void_p),
                                                         dep.ctypes.data_as(C.c_void_p), scapacity, nm.ctypes.data_as(C.c_void_p))
                        if rc != 0:
                            raise RuntimeError(L.orbx_last_error().decode())
            except Exception as e:      # noqa: re-raised after join -- a dead worker would otherwise make e2e come out too high
                errors.append(e)

        def run(nsteps):
            th = [threading.Thread(target=worker, args=(t, nsteps)) for t in range(nthreads)]
            for x in th:
                x.start()
            for x in th:
                x.join()
            if errors:
                raise errors[0]

        run(1)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run(args.steps)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    ns = len(sides)
    h2d = B * ns * ((H - 1) * P32 + W)
    d2h = B * ns * (cap * 60 + 4) + (B * (2 * 4 * scapacity + 4) if stereo else 0)
    if stereo:
        how = "orbx_extract_batch (left, right) + orbm_stereo_matches_batch on pinned host frames into pinned host results, %d handle sets x %d-pair chunks on %d host threads" % (nthreads, chunk, nthreads)
    else:
        how = "orbx_pool_submit_batch / orbx_pool_wait (the library's multi-GPU driver, this rank's GPU; two steps in flight): pinned host frames in, pinned host results out, %d lanes x %d-frame chunks" % (nthreads, chunk)
    value = B * world * args.steps / e2e_s
    ceiling = {"h2d_GBps_this_rank": h2d_gbs, "e2e_h2d_GBps_this_rank": value / world * h2d / B / 1e9,
               "frac": (value / world * h2d / B / 1e9) / h2d_gbs,
               "note": "ceiling = the same pinned frames copied host-to-device and nothing else, all ranks at once; frac = input bytes/s of the e2e leg over it"}
    return value, h2d, d2h, how, ceiling


def match_leg(local):
    import torch
    import orbx
    from orbx import synth
    match = {}
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
        # the kernel EXECUTES 6 POPC + 4 LOP3 per pair (two carry-save adders fold six xor words); algorithmically a pair is 8 POPC32
        match["%dx%d" % (N, N)] = {"Mpairs_per_s": N * N / (t_ms / 1e3) / 1e6, "ms": t_ms,
                                    "algorithmic_popc32_per_s": 8.0 * N * N / (t_ms / 1e3), "executed_popc32_per_s": 6.0 * N * N / (t_ms / 1e3)}
    return match


def finish_ours(args, cfg, S):
    """e2e leg, matcher leg, roofline / cpu_baseline objects and the JSON line (rank 0)."""
    rank, world, local, dist, B = S["rank"], S["world"], S["local"], S["dist"], S["B"]
    value, acc = S["value"], S["acc"]
    e2e_value, h2d, d2h, how, ceiling = e2e_leg(args, cfg, S)
    match = match_leg(local) if rank == 0 else {}
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
    alg = stage_alg_bytes(cfg)
    top = max(acc, key=acc.get)
    top_bytes = alg[top] if alg[top] > 0 else alg["fast_cells"]
    achieved = top_bytes * B / (acc[top] / 1e3) / 1e9
    unit_bytes = alg_bytes_per_unit(cfg)
    pipeline_gbs = unit_bytes * value / world / 1e9
    # DRAM bytes and warp-instruction counts come from an ncu capture (profiles/<tag>_traffic.json).  They are only quoted when the
    # capture was taken on THIS configuration and on the kernel sources of this checkout (stamp), otherwise the fields are null.
    traffic, traffic_src, issue = None, "no ncu capture for this configuration and these kernel sources", None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", PROFILE_TAG + "_traffic.json")))
        if tr.get("kernel_source_sha16") == kernel_source_sha16() and int(tr.get("config", 0)) == args.config:
            per_frame = sum(tr["dram_bytes_per_frame"][k] for k in top.split("+"))
            traffic, traffic_src = per_frame * B, tr["source"]
            inst = tr.get("warp_instructions_per_frame")
            if inst:    # the resource that binds this integer path: warp-instruction issue slots (148 SMs x 4 schedulers x SM clock)
                smclk = float(peaks.get("sm_max_mhz", 1965.0)) * 1e6
                issue_peak = 148 * 4 * smclk
                frames_per_s = value / world * (2 if cfg["stereo"] else 1)
                issue = {"warp_instructions_per_frame": inst, "total_per_frame": sum(inst.values()),
                         "achieved_warp_inst_per_s": sum(inst.values()) * frames_per_s, "peak_warp_inst_per_s": issue_peak,
                         "frac": sum(inst.values()) * frames_per_s / issue_peak,
                         "note": "issue-slot utilisation of the whole extraction (ncu instruction counts x measured frames/s); "
                                 "north_star names the HBM roofline, ncu shows the integer pipes bind"}
        elif tr.get("kernel_source_sha16") != kernel_source_sha16():
            traffic_src = "stale: profiles/%s_traffic.json was captured on other kernel sources (stamp %s, now %s)" % (
                PROFILE_TAG, tr.get("kernel_source_sha16"), kernel_source_sha16())
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": top, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": top_bytes * B,
                "stage_ms": acc, "stage_alg_bytes_per_frame": alg,
                "pipeline": {"alg_bytes_per_unit": unit_bytes, "achieved": pipeline_gbs, "frac": pipeline_gbs / peak,
                             "note": "whole path per GPU: B_alg x units/s / peak" + (" (stereo: two images per unit)" if cfg["stereo"] else "")},
                "issue_slots": issue}
    cpu = None
    if world == 1 and not args.no_cpu:
        n = 8 if cfg["W"] > 1500 else 16
        cpu = cpu_baseline_object(cfg, S["hostL"][:n], S["hostR"][:n] if cfg["stereo"] else None, 12.0)
    W, H = cfg["W"], cfg["H"]
    emit({
        "metric": metric_of(cfg), "value": value, "unit": unit_of(cfg), "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": S["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": workload_of(cfg), "units_per_step_per_gpu": B, "parallelism": "units sharded per GPU, no collective",
                   "l2": "inputs larger than L2 (%d MB per step)" % (B * W * H * (2 if cfg["stereo"] else 1) >> 20),
                   "keypoints_per_frame": S["nkp_mean"], "fma_mode": 1},
        "clocks": S["clk"], "gpu_launches": S["launches"], "parity": S["parity"], "parity_frames": min(args.parity_frames, B),
        "e2e": {"value": e2e_value, "unit": unit_of(cfg), "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "how": how, "numa": S["numa"], "ceiling": ceiling},
        "roofline": roofline, "cpu_baseline": cpu, "match": match,
    })
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)        # 100 x 1.7 ms: a timed region of ~0.17 s at config 0
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=0, choices=sorted(CONFIGS), help="BASELINE.json configs index (default 0: the metric's)")
    ap.add_argument("--batch", type=int, default=0, help="units per step per GPU (0 = the configuration's default)")
    ap.add_argument("--parity-frames", type=int, default=8, help="units each rank checks against the reference build before timing")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--e2e-threads", type=int, default=0, help="host threads (= lanes / handle sets) of the e2e leg; 0 = min(8, max(3, host cores per rank))")
    ap.add_argument("--e2e-wc", type=int, default=0, help="1: e2e input frames in write-combined page-locked memory")
    ap.add_argument("--e2e-chunk", type=int, default=32, help="units per chunk of the e2e leg")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    if args.config in (2, 4) and args.e2e_chunk == 32:
        args.e2e_chunk = 8 if args.config == 4 else 16           # about the same bytes per chunk as 32 VGA frames
    if args.e2e_threads <= 0:
        args.e2e_threads = min(8, max(3, (os.cpu_count() or 4) // max(int(os.environ.get("WORLD_SIZE", "1")), 1)))
    quiet_stdout()
    if args.impl == "reference":
        run_reference(args, cfg)
    else:
        run_ours(args, cfg)


if __name__ == "__main__":
    main()
