#!/usr/bin/env python
"""Key 'details' page metrics of an .ncu-rep per kernel: python scripts/ncu_details.py rep [substr]"""
import csv, subprocess, sys
rep = sys.argv[1]
sub = sys.argv[2] if len(sys.argv) > 2 else ""
out = subprocess.run(["ncu", "-i", rep, "--page", "details", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = rows[0]
ki, mi, vi, ui, ii = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit"), h.index("ID")
want = ["Duration", "DRAM Throughput", "Memory Throughput", "Compute (SM) Throughput", "Registers Per Thread",
        "Achieved Occupancy", "Theoretical Occupancy", "Issue Slots Busy", "L1/TEX Hit Rate", "L2 Hit Rate",
        "Executed Ipc Active", "Warp Cycles Per Issued Instruction", "Elapsed Cycles", "SM Frequency",
        "Mem Busy", "Max Bandwidth", "No Eligible", "Grid Size", "Block Size", "Waves Per SM"]
seen = set()
for r in rows[1:]:
    if sub and sub not in r[ki]:
        continue
    if any(r[mi].startswith(w) for w in want):
        key = (r[ii], r[mi])
        if key in seen:
            continue
        seen.add(key)
        print(r[ii], r[ki][:36], "|", r[mi], "=", r[vi], r[ui])
