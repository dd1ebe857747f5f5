#!/usr/bin/env python3
"""Average duration per kernel from an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
d = collections.defaultdict(list)
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[ui], 1e-3)
    d[r[ki].split("(")[0][:48]].append(v * scale)
for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
    print("%-50s n=%3d avg %9.1f us  total %10.1f us" % (k, len(v), sum(v) / len(v), sum(v)))
