#!/usr/bin/env python3
"""Per-source-line roll-up of an ncu report: executed warp instructions and stall samples of one kernel, attributed to
the CUDA source line (outermost non-inlined line AND innermost line) through nvdisasm's line info of the shipped cubin.

usage: ncu_lines.py report.ncu-rep kernel_name [launch_index] [--inner]
The .so must be the build that was profiled (SASS offsets are matched by position)."""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "orb_slam2v2-1_b200", "orbx", "liborbx.so")


def sass_lines(kernel):
    """[(offset, outer_line, inner_line, text)] of the kernel's SASS, from nvdisasm -g."""
    tmp = tempfile.mkdtemp()
    subprocess.check_call(["cuobjdump", "-xelf", "all", SO], cwd=tmp, stdout=subprocess.DEVNULL)
    for f in sorted(os.listdir(tmp)):
        out = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        if kernel not in out:
            continue
        res, cur_inner, cur_outer, infn = [], 0, 0, False
        for ln in out.splitlines():
            m = re.match(r"\s*\.text\.(\S+):", ln)
            if m:
                infn = kernel in m.group(1)
                continue
            if not infn:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
            if m:
                # -lineinfo carries only the innermost location: lines of headers inlined into the kernel (intrinsics,
                # sincosf_glibc.h) are attributed to the last line of the .cu file seen before them
                if m.group(1).endswith(".cu"):
                    cur_outer = cur_inner = int(m.group(2))
                else:
                    cur_inner = -int(m.group(2))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                res.append((int(m.group(1), 16), cur_outer, cur_inner, m.group(2).strip()))
        if res:
            return res
    raise SystemExit("kernel not found in any cubin")


def main():
    rep, kernel = sys.argv[1], sys.argv[2]
    args = [a for a in sys.argv[3:] if not a.startswith("--")]
    inner = "--inner" in sys.argv
    cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kernel]
    if args:
        cmd += ["--launch-skip", args[0], "--launch-count", "1"]
    txt = subprocess.run(cmd, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hi = next(i for i, r in enumerate(rows) if "Address" in r and "Source" in r)
    hdr = rows[hi]
    ci, si, ai = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Address")
    body = []
    for r in rows[hi + 1:]:
        if len(r) <= ci or not r[ai].startswith("0x"):
            if body:
                break       # next kernel instance
            continue
        body.append(r)
    base = int(body[0][ai], 16)
    sass = {o: (lo, li, t) for o, lo, li, t in sass_lines(kernel)}
    src = open(os.path.join(ROOT, "orb_slam2v2-1_b200", "csrc", "extract_kernels.cu")).read().splitlines()
    inst, smp = collections.Counter(), collections.Counter()
    for r in body:
        off = int(r[ai], 16) - base
        lo, li, _ = sass.get(off, (0, 0, ""))
        key = li if inner else lo
        inst[key] += int(r[ci] or 0)
        smp[key] += int(r[si] or 0)
    ti, ts = sum(inst.values()), sum(smp.values())
    print("kernel %s: %d warp instructions, %d samples" % (kernel, ti, ts))
    for k in sorted(inst):
        if inst[k] * 1000 < ti and smp[k] * 1000 < ts:
            continue
        line = src[k - 1].strip()[:100] if 0 < k <= len(src) else "?"
        print("%5d  inst %5.1f%%  stall %5.1f%%  %s" % (k, 100.0 * inst[k] / ti, 100.0 * smp[k] / max(ts, 1), line))


if __name__ == "__main__":
    main()
