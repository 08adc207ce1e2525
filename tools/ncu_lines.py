#!/usr/bin/env python3
"""Per-source-line stall samples: joins `ncu --page source --csv` (SASS rows) with `nvdisasm -g` line info.
usage: ncu_lines.py <source.csv> <nvdisasm function text> [top]"""
import csv, re, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]
ix = {n: h.index(n) for n in ("Address", "Warp Stall Sampling (All Samples)", "stall_long_sb", "stall_barrier", "stall_wait", "Instructions Executed")}
base = int(rows[2][0], 16)
samp = {}
for r in rows[2:]:
    try:
        samp[int(r[0], 16) - base] = tuple(int(r[ix[n]]) for n in list(ix)[1:])
    except Exception:
        pass
line_of, cur = {}, None
for l in open(sys.argv[2]):
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", l)
    if m:
        line_of[int(m.group(1), 16)] = cur
agg = defaultdict(lambda: [0, 0, 0, 0, 0])
for a, v in samp.items():
    k = line_of.get(a)
    for i, x in enumerate(v):
        agg[k][i] += x
tot = sum(v[0] for v in agg.values())
print("total samples", tot)
src = {}
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
print("%-22s %7s %6s %7s %7s %7s %9s" % ("line", "samples", "%", "long_sb", "barrier", "wait", "inst"))
for k, v in sorted(agg.items(), key=lambda t: -t[1][0])[:top]:
    text = ""
    if k:
        import glob
        for f in glob.glob("/root/repo/jn_cuclark_b200/csrc/" + k[0]):
            if f not in src: src[f] = open(f).read().split("\n")
            text = src[f][k[1] - 1].strip()[:90]
    print("%-22s %7d %6.1f %7d %7d %7d %9d  %s" % (k and f"{k[0]}:{k[1]}", v[0], 100 * v[0] / tot, v[1], v[2], v[3], v[4], text))
