#!/usr/bin/env python3
"""Opcode histogram of one kernel of liborbx.so (cuobjdump -sass): usage sass_hist.py kernel_substring"""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "orb_slam2v2-1_b200", "orbx", "liborbx.so")
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
on, c = False, collections.Counter()
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        on = sys.argv[1] in m.group(1)
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", ln)
    if on and m:
        c[m.group(1)] += 1
print(sum(c.values()), "instructions")
print("  ".join("%s:%d" % kv for kv in c.most_common(30)))
