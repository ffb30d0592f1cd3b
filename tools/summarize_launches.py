#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (ns).
usage: python tools/summarize_launches.py gpurun_out/launches.csv [--by-name] [> profiles/rNN_launches.txt]
--by-name merges the launch shapes of a kernel (NJ's scan shrinks its grid every join)."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, gi, bi = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size"), hdr.index("Block Size")
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    name = r[ki].split("(")[0].replace("<unnamed>::", "").replace("void ", "")
    key = (name, "*", r[bi]) if "--by-name" in sys.argv else (name, r[gi], r[bi])
    agg.setdefault(key, []).append(v)
tot = sum(sum(v) for v in agg.values())
print("%-46s %-16s %-14s %5s %12s %10s %10s %10s %6s" % ("kernel", "grid", "block", "n", "total_ns", "mean_ns", "min_ns", "max_ns", "share"))
for (name, g, b), v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
    print("%-46s %-16s %-14s %5d %12.0f %10.0f %10.0f %10.0f %5.1f%%" % (name[:46], g, b, len(v), sum(v), sum(v) / len(v), min(v), max(v), 100 * sum(v) / tot))
