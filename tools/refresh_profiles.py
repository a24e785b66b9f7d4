#!/usr/bin/env python
"""Copies one evidence run (gpurun_out/<tag>/, made by tools/evidence.sh) into profiles/: the bench lines, the ncu
launch list, `ncu --page details` and per-source-line totals of each captured kernel, and route_traffic.json
(DRAM bytes per launch of the route kernels, read from the captures).   python tools/refresh_profiles.py <tag>"""
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
rnd = sys.argv[2] if len(sys.argv) > 2 else "r02"   # prefix of the files written to profiles/
src, dst = os.path.join(ROOT, "gpurun_out", tag), os.path.join(ROOT, "profiles")
names = {"bench.json": "bench_compat.json", "bench_ref.json": "bench_reference_arm.json", "bench_spec.json": "bench_spec.json",
         "bench_headers.json": "bench_headers.json", "bench_decode.json": "bench_decode.json", "bench_encode.json": "bench_encode.json",
         "bench_decode_one.json": "bench_decode_one_segment.json", "bench_fanout.json": "bench_fanout.json",
         "launches.csv": "launches.csv", "pytest_gpu.log": "pytest_gpu.log"}
names = {a: rnd + "_" + b for a, b in names.items()}
for a, b in names.items():
    if os.path.exists(os.path.join(src, a)):
        shutil.copy(os.path.join(src, a), os.path.join(dst, b))


def raw_metrics(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units, val = rows[0], rows[1], rows[-1]
    m = {}
    for h, u, v in zip(hdr, units, val):
        try:
            x = float(v.replace(",", ""))
        except ValueError:
            continue
        scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}.get(u, 1.0)
        m[h] = x * scale
    return m


traffic = {}
for k in ("write", "probe", "tokenize"):
    rep = os.path.join(src, "prof_%s.ncu-rep" % k)
    if not os.path.exists(rep):
        continue
    with open(os.path.join(dst, rnd + "_k_%s_ncu_details.txt" % k), "w") as f:
        f.write(subprocess.run(["ncu", "-i", rep, "--page", "details"], capture_output=True, text=True).stdout)
    srcp = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    tmp = "/tmp/src_%s.csv" % k
    open(tmp, "w").write(srcp)
    with open(os.path.join(dst, rnd + "_k_%s_hot_lines.txt" % k), "w") as f:
        f.write(subprocess.run([sys.executable, os.path.join(dst, "ncu_lines.py"), tmp, "40"], capture_output=True, text=True).stdout)
    m = raw_metrics(rep)
    traffic["k_" + k] = {"dram_bytes_read": m.get("dram__bytes_read.sum"), "dram_bytes_write": m.get("dram__bytes_write.sum"),
                         "duration_us_under_ncu": m.get("gpu__time_duration.sum", 0) / (1e3 if m.get("gpu__time_duration.sum", 0) > 1e4 else 1),
                         "inst_executed": m.get("smsp__inst_executed.sum"), "lts_hit_rate_pct": m.get("lts__t_sector_hit_rate.pct"),
                         "l1_hit_rate_pct": m.get("l1tex__t_sector_hit_rate.pct"),
                         "warps_active_pct": m.get("sm__warps_active.avg.pct_of_peak_sustained_active"),
                         "threads_per_inst": m.get("smsp__thread_inst_executed_per_inst_executed.ratio"),
                         "registers": m.get("launch__registers_per_thread")}
if "k_write" in traffic and "k_probe" in traffic:
    bench = json.loads(open(os.path.join(src, "bench.json")).read().strip().splitlines()[-1])
    offs = 8.4e6  # k_offsets is not captured: its algorithmic bytes (cnt in, offsets out)
    total = sum(traffic[k]["dram_bytes_read"] + traffic[k]["dram_bytes_write"] for k in ("k_write", "k_probe")) + offs
    tok = traffic.pop("k_tokenize", None)
    json.dump({"what": "DRAM traffic of the route launch group (k_probe + k_offsets + k_write; k_write_rest finds an almost empty list on this workload), per batch of 2^20 keys against the 1 M-binding "
                       "table, from one `ncu --set full --clock-control none` capture of each kernel (gpurun_out/%s, summarised in "
                       "profiles/%s_k_*_ncu_details.txt)" % (tag, rnd),
               "dram_bytes_per_launch": total, "algorithmic_bytes_per_launch": bench["roofline"]["algorithmic_bytes_per_launch"],
               "kernels": dict(traffic, k_offsets={"dram_bytes_read": offs / 2, "dram_bytes_write": offs / 2, "note": "not captured: algorithmic bytes"}),
               "k_tokenize": tok,
               "reading": "traffic above the algorithmic bytes is what the three kernels hand each other through memory (counts, meta "
                          "words, list rows) plus index lines that miss L2; no list is read twice"},
              open(os.path.join(dst, "route_traffic.json"), "w"), indent=1)
    print("route group traffic %.1f MB, algorithmic %.1f MB" % (total / 1e6, bench["roofline"]["algorithmic_bytes_per_launch"] / 1e6))
for f in sorted(os.listdir(src)):
    if f.endswith(".json"):
        try:
            d = json.loads(open(os.path.join(src, f)).read().strip().splitlines()[-1])
            print("%-26s %10.1f M %s  %.3f ms/step%s" % (f, d["value"] / 1e6, d.get("unit", ""), d.get("ms_per_step", 0),
                                                        ("  e2e %.1f M" % (d["e2e"]["value"] / 1e6)) if d.get("e2e") else ""))
        except Exception as e:
            print(f, "unreadable", e)
print(open(os.path.join(src, "pytest_gpu.log")).read().strip().splitlines()[-2:])
