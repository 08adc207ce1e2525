#!/usr/bin/env python3
"""Per-source-line stall samples straight from a report: ncu -i <rep> --page source --csv --print-source cuda,sass > x.csv;
ncu_src.py x.csv [top] -- the rows whose SASS address is '-' are the per-line aggregates."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out, f, h = [], None, None
for r in rows:
    if not r: continue
    if r[0] == "File Path": f = r[1].split("/")[-1]; continue
    if r[0] == "Line No": h = r; continue
    if h and len(r) == len(h) and r[2] == "-":
        d = dict(zip(h[4:], r[4:]))
        g = lambda k: int(d.get(k, "0") or 0)
        out.append((g("Warp Stall Sampling (All Samples)"), f, r[0], r[1].strip()[:80], g("stall_long_sb"), g("stall_barrier"), g("stall_wait"), g("stall_short_sb"),
                    g("stall_mio"), g("stall_lg"), g("stall_not_selected") + g("stall_selected"), g("Instructions Executed")))
tot = sum(o[0] for o in out)
print("total samples", tot)
print("%-20s %7s %5s %7s %7s %6s %6s %6s %6s %6s %9s" % ("line", "samples", "%", "long_sb", "barrier", "wait", "shortsb", "mio", "lg", "issue", "inst"))
for o in sorted(out, reverse=True)[:top]:
    print("%-20s %7d %5.1f %7d %7d %6d %6d %6d %6d %6d %9d  %s" % (f"{o[1]}:{o[2]}", o[0], 100 * o[0] / max(tot, 1), o[4], o[5], o[6], o[7], o[8], o[9], o[10], o[11], o[3]))
