#!/usr/bin/env python3
"""Summarise `ncu --page source --csv` output: per kernel, stall-reason totals and the
top-N SASS instructions by stall samples.  usage: ncu_src_top.py src.csv [N]"""
import csv
import sys

path = sys.argv[1]
N = int(sys.argv[2]) if len(sys.argv) > 2 else 12
rows = list(csv.reader(open(path)))
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1][:90]
        hdr = rows[i + 1]
        j = i + 2
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            if len(rows[j]) == len(hdr):
                body.append(rows[j])
            j += 1
        ix = {h: k for k, h in enumerate(hdr)}
        stall_cols = [h for h in hdr if h.startswith("stall_")]
        tot = {h: 0 for h in stall_cols}
        samp = ix["# Samples"]
        for r in body:
            for h in stall_cols:
                try:
                    tot[h] += int(r[ix[h]])
                except ValueError:
                    pass
        allsum = sum(tot.values()) or 1
        print("=" * 100)
        print(name)
        print("stalls:", ", ".join(f"{h[6:]}={100*v/allsum:.1f}%" for h, v in sorted(tot.items(), key=lambda kv: -kv[1])[:8]))
        body.sort(key=lambda r: -int(r[samp] or 0))
        tots = sum(int(r[samp] or 0) for r in body) or 1
        for r in body[:N]:
            top = sorted(((h, int(r[ix[h]] or 0)) for h in stall_cols), key=lambda kv: -kv[1])[:2]
            print(f"{100*int(r[samp])/tots:5.1f}%  {r[ix['Source']][:70]:70s} {top}")
        i = j
    else:
        i += 1
