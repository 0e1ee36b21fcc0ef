#!/usr/bin/env python
"""Summaries of ncu outputs for profiles/ (run here, no GPU needed).
  python scripts/ncu_summary.py launches gpurun_out/launches.csv
  python scripts/ncu_summary.py raw gpurun_out/prof.ncu-rep [kernel-substring]
"""
import csv
import subprocess
import sys
from collections import defaultdict

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "lts__t_bytes.sum", "sm__cycles_elapsed.max", "smsp__issue_active.avg.pct_of_peak_sustained_active"]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    h = rows[0]
    ki, vi = h.index("Kernel Name"), h.index("Metric Value")
    d = defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        d[r[ki][:90]][0] += 1
        d[r[ki][:90]][1] += v
    tot = sum(v[1] for v in d.values())
    print(f"# per-kernel totals of {path} (gpu__time_duration.sum, serialised cold-cache): total {tot/1e6:.3f} ms")
    for k, v in sorted(d.items(), key=lambda x: -x[1][1]):
        print(f"{v[1]/1e6:12.3f} ms  n={v[0]:4d}  {100*v[1]/tot:5.1f}%  avg {v[1]/v[0]/1e3:10.1f} us  {k}")


def raw(path, sub=""):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[h.index("Kernel Name")]
        if sub and sub not in name:
            continue
        print("== " + name[:100])
        for k in KEYS:
            if k in h and r[h.index(k)] not in ("", "-nan", "nan"):
                print(f"   {k:75s} {r[h.index(k)]:>16s} {units[h.index(k)]}")


if __name__ == "__main__":
    (launches if sys.argv[1] == "launches" else raw)(*sys.argv[2:])
