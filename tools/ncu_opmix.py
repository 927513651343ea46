#!/usr/bin/env python3
"""Dynamic opcode mix of one kernel in an ncu report (executed warp instructions per SASS opcode):
tools/ncu_opmix.py report.ncu-rep kernel [launch_skip]"""
import collections, csv, io, re, subprocess, sys
rep, kernel = sys.argv[1], sys.argv[2]
cmd = ["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", kernel]
if len(sys.argv) > 3:
    cmd += ["--launch-skip", sys.argv[3], "--launch-count", "1"]
rows = list(csv.reader(io.StringIO(subprocess.run(cmd, capture_output=True, text=True).stdout)))
hi = next(i for i, r in enumerate(rows) if "Address" in r and "Source" in r)
h = rows[hi]
ci, si, ai, so = h.index("Instructions Executed"), h.index("# Samples"), h.index("Address"), h.index("Source")
c, st, started = collections.Counter(), collections.Counter(), False
for r in rows[hi + 1:]:
    if len(r) <= ci or not r[ai].startswith("0x"):
        if started:
            break
        continue
    started = True
    m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)?)", r[so])
    op = m.group(1) if m else "?"
    c[op] += int(r[ci] or 0)
    st[op] += int(r[si] or 0)
tot, ts = sum(c.values()), sum(st.values())
print("%s: %.1f M warp instructions" % (kernel, tot / 1e6))
for op, n in c.most_common(28):
    print("  %-16s %5.1f%%   stall samples %5.1f%%" % (op, 100.0 * n / tot, 100.0 * st[op] / max(ts, 1)))
