#!/usr/bin/env python
"""Prints the judged metrics of every kernel in an ncu report.  usage: ncu -i X.ncu-rep --page raw --csv | ncu_kernel_summary.py"""
import csv
import sys

KEEP = ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "gpu__time_duration.sum",
        "l1tex__t_sector_hit_rate.pct", "launch__block_size", "launch__grid_size", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "lts__t_sector_hit_rate.pct", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__thread_inst_executed_per_inst_executed.ratio", "sass__inst_executed_local_loads")
rows = list(csv.reader(sys.stdin))
hdr, units = rows[0], rows[1]
name_col = hdr.index("Kernel Name")
grid_col = hdr.index("Grid Size") if "Grid Size" in hdr else None
for r in rows[2:]:
    print(f"## {r[name_col]}" + (f"  (grid {r[grid_col]})" if grid_col is not None else ""))
    for i, h in enumerate(hdr):
        if h in KEEP or ("issue_stalled" in h and "per_issue_active" in h and float(r[i] or 0) >= 0.05):
            print(f"{h:95s} {units[i]:14s} {r[i]}")
