#!/usr/bin/env python3
"""Turn the captures of one GPU run into the committed summaries under profiles/:
  tools/make_profiles.py TAG "what changed"      reads gpurun_out/TAG_full.ncu-rep (ncu --set full of tools/profile_once.py 512)
                                                 and gpurun_out/TAG_launches_b512.csv (ncu launch list of bench.py --steps 2)
  -> profiles/TAG_ncu_full_summary.txt, profiles/TAG_traffic.json (DRAM bytes and warp instructions per frame and kernel),
     profiles/TAG_launches_b512.csv"""
import collections, csv, hashlib, io, json, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def kernel_source_sha16():
    """The stamp bench.py compares (same files, same order): counters are only quoted for the kernel sources they were taken on."""
    h = hashlib.sha256()
    for f in ("extract_kernels.cu", "extract_kernels.cuh", "orbx_extractor.cu"):
        h.update(open(os.path.join(ROOT, "orb_slam2v2-1_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]



tag, note = sys.argv[1], sys.argv[2]
rep = os.path.join(ROOT, "gpurun_out", tag + "_full.ncu-rep")
table = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "ncu_table.py"), rep], capture_output=True, text=True).stdout
open(os.path.join(ROOT, "profiles", tag + "_ncu_full_summary.txt"), "w").write(
    "# ncu --set full --clock-control none, one extraction of 512 synthetic 640x480 frames (tools/profile_once.py 512), %s\n%s" % (note, table))
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv", "--metrics", "dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum"],
                     capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw)))
h = r[0]
ki, ri, wi, ii = h.index("Kernel Name"), h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("smsp__inst_executed.sum")
mul = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
names = {"k_resize": "pyramid", "k_fast_cells": "fast_cells", "k_octree": "octree", "k_blur": "blur", "k_orient_describe": "orient_describe"}
acc, inst = collections.defaultdict(float), collections.defaultdict(float)
for row in r[2:]:
    k = [v for kk, v in names.items() if kk in row[ki]][0]
    acc[k] += float(row[ri].replace(",", "")) * mul[r[1][ri]] + float(row[wi].replace(",", "")) * mul[r[1][wi]]
    inst[k] += float(row[ii].replace(",", ""))
json.dump({"source": "profiles/%s_ncu_full_summary.txt (ncu --set full --clock-control none, tools/profile_once.py 512)" % tag,
           "kernel_source_sha16": kernel_source_sha16(), "config": 0,
           "frames_per_launch": 512, "shape": "640x480 nf=1000",
           "dram_bytes_per_frame": {k: int(round(v / 512)) for k, v in acc.items()},
           "warp_instructions_per_frame": {k: int(round(v / 512)) for k, v in inst.items()}},
          open(os.path.join(ROOT, "profiles", tag + "_traffic.json"), "w"), indent=1)
rows = list(csv.reader(open(os.path.join(ROOT, "gpurun_out", tag + "_launches_b512.csv"))))
hdr = [i for i, x in enumerate(rows) if x and x[0] == "ID"][0]
h = rows[hdr]
ki, vi, gi, bi = h.index("Kernel Name"), h.index("Metric Value"), h.index("Grid Size"), h.index("Block Size")
out = ["# ncu --metrics gpu__time_duration.sum --clock-control none -c 400: python bench.py --steps 2 --warmup 1 --no-cpu "
       "(512 frames per step as two halves of 256 on two streams; per-launch times are cold-cache and serialised)", "id,kernel,grid,block,ns"]
share = collections.defaultdict(float)
for n, x in enumerate(rows[hdr + 1:]):
    if len(x) <= vi:
        continue
    name = re.sub(r"^void ", "", x[ki]).split("(")[0].replace("orbx::", "")
    ns = int(float(x[vi].replace(",", "")))
    out.append("%d,%s,(%s),(%s),%d" % (n, name, x[gi].strip("()").replace(", ", "x"), x[bi].strip("()").replace(", ", "x"), ns))
    share[name.split("<")[0].replace("k_orient_describe32", "k_orient_describe")] += ns
open(os.path.join(ROOT, "profiles", tag + "_launches_b512.csv"), "w").write("\n".join(out) + "\n")
tot = sum(v for k, v in share.items() if k in names)
print(table)
print("share of the step (launch list):", {names[k]: round(v / tot, 3) for k, v in share.items() if k in names})
print("warp instructions per frame:", dict(inst and {k: int(v / 512) for k, v in inst.items()}))
