#!/usr/bin/env python3
"""Static opcode histogram of the kernels of liborbx.so (cuobjdump -sass), one block per kernel whose mangled name contains one of
the given substrings (default: every extraction and matcher kernel).  usage: sass_hist.py [substring ...] > profiles/rNN_sass_hist.txt
Static counts (instructions in the binary, not executed): what they prove is WHICH instructions a kernel is made of -- UTMALDG /
SYNCS (TMA + mbarrier), LDGSTS, IDP, VIMNMX3, VABSDIFF4, POPC, REDUX ..."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
so = os.path.join(ROOT, "orb_slam2v2-1_b200", "orbx", "liborbx.so")
want = sys.argv[1:] or ["k_resize", "k_fast_cells", "k_octree", "k_blur", "k_orient_describe", "k_bf_scan", "k_stereo_batch", "k_area_best"]
out = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
def demangle(n):
    d = subprocess.run(["cu++filt", n], capture_output=True, text=True).stdout.strip() or n
    m = re.match(r"(.*?>|[^(<]*)\((?!int\)|bool\))", d)       # drop the parameter list, keep template arguments such as <(int)64, (bool)1>
    return m.group(1) if m else d
NOTABLE = ("UTMALDG", "SYNCS", "LDGSTS", "IDP", "VIMNMX3", "VIMNMX", "VABSDIFF4", "POPC", "REDUX", "SHFL", "VOTE", "BAR", "PRMT", "LDS", "LDG", "STG", "ATOMS")
cur, hist = None, collections.OrderedDict()
for ln in out.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        cur = m.group(1) if any(w in m.group(1) for w in want) else None
        if cur:
            hist[cur] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+(?:\.[A-Z0-9_]+)*)", ln)
    if cur and m:
        op = m.group(1)
        base = op.split(".")[0]
        # keep the qualifiers that matter for the claims in DESIGN.md, fold the rest
        key = op if base in ("UTMALDG", "SYNCS", "LDGSTS", "IDP", "VIMNMX3", "VIMNMX", "REDUX", "VABSDIFF4") else base
        key = ".".join(key.split(".")[:3])
        hist[cur][key] += 1
print("liborbx.so sha256[:16] =", __import__("hashlib").sha256(open(so, "rb").read()).hexdigest()[:16])
for name, c in hist.items():
    print("\n%s   (%d instructions)" % (demangle(name), sum(c.values())))
    print("  " + "  ".join("%s:%d" % kv for kv in c.most_common(40)))
    print("  notable: " + "  ".join("%s:%d" % (k, v) for k, v in sorted(c.items()) if k.split(".")[0] in NOTABLE))
