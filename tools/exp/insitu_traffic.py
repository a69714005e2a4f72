#!/usr/bin/env python3
"""Summarise an `ncu --cache-control none --metrics gpu__time_duration.sum,dram__bytes_read.sum,
dram__bytes_write.sum --csv` launch list: per kernel name, mean us / MB read / MB written over
the last N steps (kernels run with the cache state the previous kernel left: in-situ traffic)."""
import csv
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, mi, vi, ii = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
d = OrderedDict()
for r in rows[1:]:
    d.setdefault(int(r[ii]), {"name": r[ki].split("(")[0][-60:]})[r[mi]] = float(r[vi].replace(",", ""))
ids = sorted(d)
ids = ids[len(ids) // 2:]  # steady state
agg = OrderedDict()
for i in ids:
    m = d[i]
    a = agg.setdefault(m["name"], [0, 0.0, 0.0, 0.0])
    a[0] += 1
    a[1] += m.get("gpu__time_duration.sum", 0) / 1e3
    a[2] += m.get("dram__bytes_read.sum", 0) / 1e6
    a[3] += m.get("dram__bytes_write.sum", 0) / 1e6
tot = [0.0, 0.0, 0.0]
for k, a in agg.items():
    print(f"{k:62s} n={a[0]:3d} us={a[1]/a[0]:8.1f} readMB={a[2]/a[0]:8.1f} writeMB={a[3]/a[0]:8.1f}")
