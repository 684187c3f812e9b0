#!/usr/bin/env python
"""Summarise an ncu report's source page: stall samples and executed instructions per
region of N SASS instructions.  usage: ncu_regions.py report.ncu-rep [chunk]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 250
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, data = rows[1], rows[2:]
ci = {n: i for i, n in enumerate(h)}
S, I, SRC = ci["# Samples"], ci["Instructions Executed"], ci["Source"]
stall = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
tot = sum(int(r[S] or 0) for r in data)
tot_i = sum(int(r[I] or 0) for r in data)
print(rows[0][1], "| samples", tot, "| warp instructions", tot_i, "| static", len(data))
agg = {n: sum(int(r[ci[n]] or 0) for r in data) for n in stall}
print("stalls:", [(k[6:], round(100 * v / tot, 1)) for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]])
for c0 in range(0, len(data), chunk):
    ch = data[c0:c0 + chunk]
    s = sum(int(r[S] or 0) for r in ch)
    ins = sum(int(r[I] or 0) for r in ch)
    if ins == 0:
        continue
    ops = collections.Counter((r[SRC].split()[1] if r[SRC].startswith("@") else r[SRC].split()[0]).split(".")[0] for r in ch)
    st = {n: sum(int(r[ci[n]] or 0) for r in ch) for n in stall}
    top = sorted(st.items(), key=lambda x: -x[1])[:3]
    print("%5d samp %5.1f%% inst %5.1f%% %s %s" % (c0, 100 * s / tot, 100 * ins / tot_i, ops.most_common(3),
                                                  [(k[6:], round(100 * v / max(s, 1))) for k, v in top]))
