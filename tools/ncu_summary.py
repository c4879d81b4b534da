#!/usr/bin/env python3
"""Key metrics per kernel from an .ncu-rep (run where ncu is installed; no GPU needed):
    python tools/ncu_summary.py gpurun_out/x.ncu-rep > profiles/x_summary.txt"""
import csv
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__t_sector_hit_rate.pct",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "launch__grid_size", "launch__block_size"]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    name = hdr.index("Kernel Name")
    groups = {}
    for r in data:
        groups.setdefault(r[name], []).append(r)
    print("# %s: median launch of each kernel (ncu --set full, cold L2, clocks not locked)" % rep.split("/")[-1])
    for k, v in groups.items():
        print("\n%s   [%d launches captured]" % (k, len(v)))
        col = hdr.index("gpu__time_duration.sum")
        v.sort(key=lambda r: float(r[col].replace(",", "")))
        r = v[len(v) // 2]
        for w in WANT:
            if w in hdr:
                i = hdr.index(w)
                print("  %-72s %14s %s" % (w, r[i], units[i]))


if __name__ == "__main__":
    main()
