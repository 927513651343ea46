#!/usr/bin/env python3
"""Readable per-kernel table of an ncu report: tools/ncu_table.py report.ncu-rep"""
import csv, io, subprocess, sys
M = [("gpu__time_duration.sum", "us"), ("launch__registers_per_thread", "regs"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
     ("smsp__inst_executed.sum", "Minst"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
     ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"), ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
     ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"), ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1%"),
     ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
     ("dram__bytes_read.sum", "rdMB"), ("dram__bytes_write.sum", "wrMB"),
     ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "st_long"),
     ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "st_short"),
     ("smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "st_math"),
     ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "st_wait"),
     ("smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "st_bar"),
     ("smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "st_mio"),
     ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "st_lg")]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv", "--metrics", ",".join(m for m, _ in M)], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, units = rows[0], rows[1]
print("kernel".ljust(20) + " ".join(n.rjust(8) for _, n in M))
for r in rows[2:]:
    name = r[h.index("Kernel Name")].split("(")[0].replace("orbx::", "")
    vals = []
    for m, n in M:
        v = float(r[h.index(m)].replace(",", "")) if m in h and r[h.index(m)] not in ("", "n/a") else float("nan")
        u = units[h.index(m)] if m in h else ""
        if n == "us":
            v *= {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}.get(u, 1)
        if n == "Minst":
            v /= 1e6
        if n in ("rdMB", "wrMB"):
            v *= {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1, "Gbyte": 1e3}.get(u, 1)
        vals.append(("%.1f" % v).rjust(8))
    print(name[:20].ljust(20) + " ".join(vals))
