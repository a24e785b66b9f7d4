#!/usr/bin/env python
"""Per-source-line instruction and stall totals of an ncu report (see ncu_lines.py), with file names:
   ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv; python ncu_regions.py src.csv [top]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
hdr, agg, stall = None, {}, {}
for r in rows:
    if r and r[0] == "Line No":
        hdr = r
        ii, si = hdr.index("Instructions Executed"), hdr.index("Warp Stall Sampling (All Samples)")
        continue
    if hdr is None or len(r) <= ii or r[0] == "":
        continue
    try:
        key = (int(r[0]), r[1].strip())
        agg[key] = agg.get(key, 0) + int(r[ii] or 0)
        stall[key] = stall.get(key, 0) + int(r[si] or 0)
    except ValueError:
        pass
tot, ts = sum(agg.values()) or 1, sum(stall.values()) or 1
print("total warp instructions %.1fM, stall samples %d" % (tot / 1e6, ts))
for key, v in sorted(agg.items(), key=lambda x: -x[1])[:top]:
    print("%5.1f%% inst %5.1f%% stall  L%-4d %s" % (100 * v / tot, 100 * stall[key] / ts, key[0], key[1][:105]))
