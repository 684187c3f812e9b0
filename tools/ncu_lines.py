#!/usr/bin/env python
"""Stall samples of an ncu report attributed to CUDA source lines (needs -lineinfo + --import-source on).
usage: python tools/ncu_lines.py report.ncu-rep [top_n]"""
import csv
import os
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
fname, lines, S, I = "?", [], None, None
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        fname = os.path.basename(r[1])
    elif r[0] == "Line No":
        S, I = r.index("# Samples"), r.index("Instructions Executed")
    elif r[0].isdigit() and S is not None and len(r) > I:
        try:        # a source line that itself holds quotes and commas splits wrongly: skip it
            lines.append((int(r[S] or 0), int(r[I] or 0), fname, int(r[0]), r[1].strip()))
        except ValueError:
            pass
tot = sum(x[0] for x in lines) or 1
tot_i = sum(x[1] for x in lines) or 1
print("samples", tot, "warp instructions", tot_i)
for s, i, f, n, src in sorted(lines, key=lambda x: -x[0])[:top_n]:
    print("%5.1f%% samp %5.1f%% inst  %s:%-4d %s" % (100 * s / tot, 100 * i / tot_i, f, n, src[:105]))
