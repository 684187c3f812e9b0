#!/usr/bin/env python
"""Aggregate an `ncu --csv --metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum]` launch list
per kernel: launches, total time, DRAM bytes.  usage: python tools/ncu_launches.py launches.csv [divide_by]"""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
div = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
hdr, agg = None, collections.OrderedDict()
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
for r in rows:
    if r[0] == "ID":
        hdr = r
        continue
    if hdr is None:
        continue
    d = dict(zip(hdr, r))
    a = agg.setdefault(d["Kernel Name"][:64], {"n": 0, "t": 0.0, "rd": 0.0, "wr": 0.0})
    v, unit, m = float(d["Metric Value"].replace(",", "")), d["Metric Unit"], d["Metric Name"]
    if m == "gpu__time_duration.sum":
        a["n"] += 1
        a["t"] += {"nsecond": 1e-3, "ns": 1e-3, "usecond": 1.0, "us": 1.0, "msecond": 1e3, "ms": 1e3}.get(unit, 1.0) * v
    elif m == "dram__bytes_read.sum":
        a["rd"] += v * scale.get(unit, 1)
    elif m == "dram__bytes_write.sum":
        a["wr"] += v * scale.get(unit, 1)
tot = sum(a["t"] for a in agg.values())
for k, a in agg.items():
    print("%-66s n=%4d  t=%10.1f us (%5.1f%%)  rd=%9.1f MB  wr=%9.1f MB" % (k, a["n"], a["t"] / div, 100 * a["t"] / tot, a["rd"] / div / 1e6, a["wr"] / div / 1e6))
print("total %.1f us" % (tot / div))
