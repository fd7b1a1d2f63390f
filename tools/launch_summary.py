#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = None
agg = collections.OrderedDict()
for r in rows:
    if "Kernel Name" in r:
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        key = d["Kernel Name"].split("(")[0][:48] + " grid " + d["Grid Size"] + " block " + d["Block Size"]
        agg.setdefault(key, []).append(float(d["Metric Value"].replace(",", "")))
tot = sum(sum(v) for v in agg.values())
for k, v in agg.items():
    print(f"{k:90s} n={len(v):4d} mean {sum(v)/len(v)/1e3:9.2f} us  share {100*sum(v)/tot:5.1f}%")
