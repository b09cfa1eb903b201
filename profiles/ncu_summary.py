#!/usr/bin/env python
"""Summarise an .ncu-rep (--set full) per launch: duration, DRAM bytes, throughput %, occupancy, top stall reasons.
usage: ncu_summary.py file.ncu-rep"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "smsp__inst_executed.sum", "sm__inst_executed_pipe_lsu.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print("==", d["Kernel Name"][:110])
        for k in KEYS:
            if k in d:
                print(f"   {k:70s} {d[k]:>16s} {units[hdr.index(k)]}")
        stalls = [(float(d[k].replace(',', '')), k) for k in hdr if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("per_issue_active.ratio") and d[k]]
        if not stalls:
            stalls = [(float(d[k].replace(',', '')), k) for k in hdr if "issue_stalled" in k and k.endswith(".pct") and d[k]]
        for v, k in sorted(stalls, reverse=True)[:7]:
            print(f"   stall {k:64s} {v:10.3f}")


if __name__ == "__main__":
    main()
