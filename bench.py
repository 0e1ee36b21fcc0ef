#!/usr/bin/env python
"""bench.py -- headline benchmark of the compressed-genotype hot path on B200.

Metric (BASELINE.json): dgemm_compressed throughput in SNP*indiv*col/s on configs[2]
(1M SNPs x 100k individuals, ncol=10, 'n' and 't'), sharded by SNP blocks over N GPUs
("strong" scaling: the total matrix is fixed; 't' needs no exchange, 'n' is followed by one NCCL
all-reduce of the indiv x ncol partials).  A "step" = one 'n' call + one 't' call.

  value     device-resident B/C (mb200_dgemm_compressed_device), CUDA events on the library's stream
  e2e       the reference-facing C-ABI call dgemm_compressed() with pinned HOST B/C, copies inside
  roofline  dominant kernel's algorithmic HBM bytes / its CUDA-event duration vs MEASURED_PEAKS.json
  cpu_baseline  the unmodified reference CPU 5codes library (oracle/_ref) on a bounded sub-block
  grm       the GRM crossproduct of configs[1] (50k x 10k) in int8 TOPS (second half of the metric)

`--impl reference` times the reference CPU library alone (rank 0 only).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "dgemm_compressed SNP*indiv*col/s"
UNIT = "SNP*indiv*col/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--snps", type=int, default=1_000_000)
    ap.add_argument("--indiv", type=int, default=100_000)
    ap.add_argument("--ncol", type=int, default=10)
    ap.add_argument("--no-grm", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--grm-snps", type=int, default=50_000)
    ap.add_argument("--grm-indiv", type=int, default=10_000)
    ap.add_argument("--cpu-snps", type=int, default=100_000)
    ap.add_argument("--cpu-indiv", type=int, default=10_000)
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops", 1590.0)), "measured"
    return 6650.0, 1590.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4)
                          if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# reference CPU arm (oracle/_ref = the unmodified reference CPU 5codes library)
# ------------------------------------------------------------------------------------------------
def cpu_reference_run(plink, snps, indiv, f, ncol, steps, warmup):
    """Time `steps` steps ('N' + 'T') of the reference CPU library on the given sample."""
    import numpy as np
    from oracle import ref
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)
    ref.set_options(cores=cores, centered=True, variant=256)
    h = ref.RefCompressed(plink, snps, indiv, f, ncol)
    rng = np.random.default_rng(0)
    Bn = np.asfortranarray(rng.standard_normal((snps, ncol)))
    Bt = np.asfortranarray(rng.standard_normal((indiv, ncol)))
    Cn = np.zeros((indiv, ncol), order="F")
    Ct = np.zeros((snps, ncol), order="F")
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        h.dgemm("N", Bn, Cn)
        h.dgemm("T", Bt, Ct)
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    units = 2.0 * snps * indiv * ncol
    return units / (sum(times) / len(times)), units / min(times), cores, sum(times) / len(times)


def host_sample(snps, indiv, seed):
    """Sample bed with the benchmark's distribution (numpy twin of the device simulator's law)."""
    from oracle import simulate
    plink, _, f = simulate.simulate(snps, indiv, seed=seed)
    return plink, f


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref
    if not ref.available() and not ref.build():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_miraculix.so not built"}))
        return
    snps, indiv = args.cpu_snps, args.cpu_indiv
    plink, f = host_sample(snps, indiv, 1)
    mean_v, best_v, cores, sec = cpu_reference_run(plink, snps, indiv, f, args.ncol, args.steps,
                                                   args.warmup)
    sample = (f"{snps} SNPs x {indiv} individuals sub-block of the workload, ncol={args.ncol}, "
              f"'N'+'T' per step, reference CPU 5codes (variant 256), all host threads")
    line = {"metric": METRIC, "value": mean_v, "unit": UNIT, "impl": "reference", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": workload_name(args), "sample": sample},
            "cpu_baseline": {"value": mean_v, "unit": UNIT, "cores": cores, "kind": "reference",
                             "sample": sample},
            "e2e": {"value": mean_v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


def workload_name(args):
    return (f"dgemm_compressed 'n'+'t', ncol={args.ncol}, {args.snps} SNPs x {args.indiv} "
            f"individuals (BASELINE configs[2]), SNP-block sharded over {args.gpus} GPU(s)")
