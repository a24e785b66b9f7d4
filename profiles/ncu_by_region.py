#!/usr/bin/env python
"""Instruction totals of an ncu source page grouped by function region of this repo's kernels.
   python ncu_by_region.py src.csv   (src.csv from: ncu -i X --page source --csv --print-source cuda,sass)"""
import csv
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CS = os.path.join(ROOT, "chanamq_b200", "csrc")
regions = {}  # stripped source line -> region name
for fn in ("chq_common.h", "probe.cuh", "merge.cuh", "weave.cuh", "flat.cuh", "kernels.cu"):
    region = fn
    for ln, line in enumerate(open(os.path.join(CS, fn)), 1):
        m = re.match(r"^(?:template.*\n)?(?:__device__|__global__|CHQ_HD).*?(\w+)\(", line)
        if m:
            region = fn + ":" + m.group(1)
        if fn == "kernels.cu":
            for tag in ("phase A", "tile prefix sum", "phase B"):
                if "---- " + tag in line:
                    region = "kernels.cu:k_route " + tag
        regions.setdefault((ln, line.strip()), region)
rows = list(csv.reader(open(sys.argv[1])))
hdr, tot = None, {}
for r in rows:
    if r and r[0] == "Line No":
        hdr = r
        ii = hdr.index("Instructions Executed")
        continue
    if hdr is None or len(r) <= ii or r[0] == "":
        continue
    try:
        key = (int(r[0]), r[1].strip())
        v = int(r[ii] or 0)
    except ValueError:
        continue
    reg = regions.get(key, "other (intrinsics headers)")
    tot[reg] = tot.get(reg, 0) + v
s = sum(tot.values()) or 1
for k, v in sorted(tot.items(), key=lambda x: -x[1]):
    print("%6.1fM %5.1f%%  %s" % (v / 1e6, 100 * v / s, k))
print("%6.1fM total" % (s / 1e6))
