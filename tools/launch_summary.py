#!/usr/bin/env python3
"""Condense an `ncu --metrics gpu__time_duration.sum --csv` launch list into time per kernel.

    python tools/launch_summary.py gpurun_out/launches.csv > profiles/rN_launch_summary.csv"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 10]
hdr = rows[0]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
tot = collections.defaultdict(float); cnt = collections.Counter()
for r in rows[1:]:
    name = re.sub(r"\(.*$", "", r[ki]).replace("void ", "").strip()
    v = float(r[vi].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
    tot[name] += v; cnt[name] += 1
total = sum(tot.values())
w = csv.writer(sys.stdout)
w.writerow(["kernel", "launches", "total_ms", "share_of_gpu_time"])
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    w.writerow([k, cnt[k], round(v, 3), round(v / total, 4)])
