#!/usr/bin/env python3
"""Text summary of one .ncu-rep (ncu --set full): the metrics DESIGN.md and bench.py cite.
usage: ncu_summary.py report.ncu-rep > profiles/NAME.txt"""
import csv, io, subprocess, sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__issue_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
    "sm__icc_request_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
]
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    d = dict(zip(hdr, r)); u = dict(zip(hdr, units))
    print("kernel:", d.get("Kernel Name"), " id:", d.get("ID"))
    for k in KEYS:
        if k in d: print("  %-58s %s %s" % (k, d[k], u.get(k, "")))
    print("  warp stall reasons (warps stalled per issue-active cycle):")
    st = [(float(d[k]), k) for k in hdr if "average_warps_issue_stalled" in k and k.endswith("per_issue_active.ratio")]
    for v, k in sorted(st, reverse=True)[:8]:
        print("    %-28s %.3f" % (k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""), v))
