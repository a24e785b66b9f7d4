#!/usr/bin/env python
"""Aggregate an ncu report's source page per CUDA source line:
   ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv ; python ncu_lines.py src.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None
agg = []
files = None
for r in rows:
    if r and r[0] == "File Name":
        files = r[1] if len(r) > 1 else None
    if r and r[0] == "Line No":
        hdr = r
        ii = hdr.index("Instructions Executed")
        si = hdr.index("Warp Stall Sampling (All Samples)")
        continue
    if hdr is None or len(r) <= ii:
        continue
    if r[0] != "":  # a source line header row
        try:
            agg.append([int(r[0]), r[1].strip(), int(r[ii] or 0), int(r[si] or 0), files])
        except ValueError:
            pass
tot_i = sum(a[2] for a in agg) or 1
tot_s = sum(a[3] for a in agg) or 1
print("total warp-instructions %d, stall samples %d" % (tot_i, tot_s))
print("--- by instructions executed")
for a in sorted(agg, key=lambda a: -a[2])[:top]:
    print("%5.1f%% inst %5.1f%% stall  L%-4d %s" % (100 * a[2] / tot_i, 100 * a[3] / tot_s, a[0], a[1][:110]))
print("--- by stall samples")
for a in sorted(agg, key=lambda a: -a[3])[:top]:
    print("%5.1f%% inst %5.1f%% stall  L%-4d %s" % (100 * a[2] / tot_i, 100 * a[3] / tot_s, a[0], a[1][:110]))
