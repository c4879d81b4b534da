#!/usr/bin/env python3
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per (kernel, grid) the number
of launches, the median duration and the share of all GPU kernel time.
    python tools/launch_summary.py gpurun_out/launches.csv "title line" > profiles/...summary.txt"""
import csv
import statistics
import sys

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r.get("Metric Unit", "us")
    if unit.startswith("ns"):      # durations as the first column unit (ns / us / ms)
        v /= 1e3
    elif unit.startswith("ms"):
        v *= 1e3
    rows.append((r["Kernel Name"], r.get("Grid Size", ""), v))
groups = {}
for name, grid, v in rows:
    groups.setdefault((name, grid), []).append(v)
total = sum(v for _, _, v in rows)
print("# " + (sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]))
print("# kernel, grid, launches, median us, total us, share of all GPU kernel time (cold-cache, serialised: compare shares)")
for (name, grid), vs in sorted(groups.items(), key=lambda kv: -sum(kv[1])):
    print("%-62s %-16s %5d %9.2f %10.1f %6.1f%%" % (name[:62], grid, len(vs), statistics.median(vs), sum(vs), 100 * sum(vs) / total))
