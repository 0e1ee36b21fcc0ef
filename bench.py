#!/usr/bin/env python
"""bench.py -- headline benchmark of the compressed-genotype hot path on B200.

Metric (BASELINE.json): dgemm_compressed throughput in SNP*indiv*col/s on configs[2]
(1M SNPs x 100k individuals, ncol=10, 'n' and 't'), sharded by SNP blocks over N GPUs
("strong" scaling: the total matrix is fixed; 't' needs no exchange, 'n' is followed by one NCCL
all-reduce of the indiv x ncol partials).  A "step" = one 'n' call + one 't' call.

  value     device-resident B/C (mb200_dgemm_compressed_device), CUDA events on the library's stream;
            N > 1: one process per GPU (torch.distributed / NCCL all-reduce of the 'n' partials)
  e2e       the reference-facing C-ABI call dgemm_compressed() with HOST B/C, copies inside; N > 1:
            ONE process (rank 0) drives all N GPUs through the library's single-process multi-device
            handle (MIRACULIX_NGPUS path, csrc/multi.cu) -- what an unchanged Julia / Fortran caller
            gets; pinned host panels for the headline, pageable ones beside it (e2e.pageable)
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
    ap.add_argument("--no-pcg", action="store_true")
    ap.add_argument("--pcg-snps", type=int, default=500_000)
    ap.add_argument("--pcg-indiv", type=int, default=200_000)
    ap.add_argument("--pcg-iters", type=int, default=20)
    ap.add_argument("--grm-snps", type=int, default=50_000)
    ap.add_argument("--grm-indiv", type=int, default=10_000)
    ap.add_argument("--grm-full", type=int, default=-1,
                    help="configs[4] (1M SNPs x 100k individuals GRM over the GPUs): 1 on, 0 off, -1 = on at 8 GPUs")
    ap.add_argument("--grm-full-snps", type=int, default=1_000_000)
    ap.add_argument("--grm-full-indiv", type=int, default=100_000)
    ap.add_argument("--no-small", action="store_true", help="skip the configs[0] latency leg")
    ap.add_argument("--wide-ncol", type=int, default=50,
                    help="extra leg: 'n'+'t' with this many columns (gblup PCA shape); 0 = off")
    ap.add_argument("--e2e-devices", default="",
                    help="device list of the single-process e2e arm (default 0..N-1; '0,0' exercises the "
                         "multi-device handle on one GPU)")
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
             "clocks_event_reasons.sw_power_cap,power.draw")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def mark(self):
        """Number of samples taken so far (to split the record into phases)."""
        return len(self.rows)

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        return self.stats(self.rows)

    @staticmethod
    def stats(rows):
        sm = [float(r[0]) for r in rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in rows if len(r) >= 6 for i in range(4)
                          if r[2 + i].lower().startswith("active")})
        pw = []
        for r in rows:
            try:
                pw.append(float(r[6]))
            except (IndexError, ValueError):
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm),
                "power_w": statistics.median(pw) if pw else None}


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


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def dgemm_bytes(rows, k, ncol):
    """ALGORITHMIC HBM bytes of one dgemm_compressed call: packed matrix of the
    orientation used, read once, unpadded + B + C."""
    return rows * ((k + 3) // 4) + 8 * ncol * k + 8 * ncol * rows


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from miraculix_b200 import api, build
    from miraculix_b200.sharded import ShardedCompressed

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this product has no CPU path "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        # host-side barriers for the phases in which ONE rank drives all GPUs (a NCCL barrier would
        # park a spinning kernel on every waiting rank's GPU)
        host_group = dist.new_group(backend="gloo")
    build.build()
    lib = api.load_library()
    api.set_options(use_gpu=True, not_center=False)

    snps, indiv, ncol = args.snps, args.indiv, args.ncol
    sh = ShardedCompressed(None, snps, indiv, rank, world)
    s_loc = sh.s_count
    bs = (indiv + 3) // 4

    # synthetic shard, generated on the device (never crosses PCIe), then repacked on the device
    t0 = time.perf_counter()
    bed = torch.empty(s_loc * bs, dtype=torch.uint8, device=dev)
    SEED = 1000
    api._check(lib.mb200_simulate_plink_rows_device(C.c_void_p(bed.data_ptr()), s_loc, indiv, SEED,
                                                    sh.s_begin), "simulate")
    handle = C.c_void_p(None)
    api._check(lib.mb200_plink2compressed_device(C.c_void_p(bed.data_ptr()), s_loc, indiv, None,
                                                 ncol, C.byref(handle)), "plink2compressed_device")
    torch.cuda.synchronize()
    del bed
    torch.cuda.empty_cache()
    setup_s = time.perf_counter() - t0
    sh.handle = handle
    stream = sh.stream()

    # B panels: generated once on the host (seeded), so that the device-resident arm, the sharded
    # arm and the single-process multi-device arm multiply the same numbers
    rng = np.random.default_rng(7)
    hBn_full = torch.from_numpy(rng.standard_normal((ncol, snps))).pin_memory()
    hBt = torch.from_numpy(rng.standard_normal((ncol, indiv))).pin_memory()
    Bn = hBn_full[:, sh.s_begin:sh.s_begin + s_loc].contiguous().to(dev)   # this rank's rows of B
    Bt = hBt.to(dev)
    Cn = torch.empty((ncol, indiv), dtype=torch.float64, device=dev)
    Ct = torch.empty((ncol, s_loc), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    def step_device():
        sh.dgemm_n(Bn, Cn)
        sh.dgemm_t(Bt, Ct)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        launches0 = api.kernel_launches()
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), api.kernel_launches() - launches0

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    mark0 = sampler.mark() if sampler else 0
    total_ms, launches = timed(step_device, args.steps, args.warmup)
    mark1 = sampler.mark() if sampler else 0
    # per-launch durations of the dominant kernel inside the timed region (event ring)
    nlast = min(2 * args.steps, 128)
    ms_arr = (C.c_float * nlast)()
    tr_arr = (C.c_int * nlast)()
    got = lib.mb200_dgemm_kernel_times(handle, nlast, ms_arr, tr_arr)
    k_n = [ms_arr[i] for i in range(got) if tr_arr[i] == 0]
    k_t = [ms_arr[i] for i in range(got) if tr_arr[i] == 1]

    # results of the device-timed arm, kept for the parity check of the e2e arms below
    stream.synchronize()
    Cn_dev, Ct_dev = Cn.cpu(), Ct.cpu()

    clocks = None
    e2e_devices = [int(x) for x in args.e2e_devices.split(",")] if args.e2e_devices else list(range(world))
    e2e_steps = max(2, min(args.steps, 5))

    def host_barrier():
        if host_group is not None:
            dist.barrier(group=host_group)

    # ---- e2e: the reference-facing call with HOST panels, copies inside the timed region.  One
    # process: at N > 1 rank 0 alone drives all GPUs through a sharded handle of the SAME matrix
    # (the other ranks keep their contexts but are idle on a host-side barrier).
    torch.cuda.synchronize()
    host_barrier()
    e2e = None
    if rank == 0:
        e2e = run_e2e(args, lib, api, handle if len(e2e_devices) == 1 else None, e2e_devices, SEED,
                      hBn_full, hBt, Cn_dev, Ct_dev, e2e_steps)
    host_barrier()
    if sampler:
        sampler.stop()
        clocks = ClockSampler.stats(sampler.rows)               # whole run: device-timed + e2e legs
        clocks["device_timed"] = ClockSampler.stats(sampler.rows[mark0:mark1])
    small = None
    if rank == 0 and not args.no_small:
        small = run_small(lib, api)
    wide = None
    if rank == 0 and world == 1 and args.wide_ncol > 0 and not args.no_small:
        try:
            wide = run_wide(args, lib, api, dev)
        except Exception as exc:
            wide = {"error": repr(exc)}
    host_barrier()
    pcg = None
    if not args.no_pcg:
        pcg = run_pcg(args, lib, api, dev, rank, world, host_barrier, e2e_devices)
    grm_part = None
    if not args.no_grm and world > 1:
        grm_part = run_grm_partition(args, lib, api, dev, rank, world, args.grm_snps, args.grm_indiv,
                                     "BASELINE configs[1]", host_barrier=host_barrier, e2e_devices=e2e_devices)
    grm_full = None
    if not args.no_grm and world > 1 and (args.grm_full == 1 or (args.grm_full < 0 and world == 8)):
        grm_full = run_grm_partition(args, lib, api, dev, rank, world, args.grm_full_snps,
                                     args.grm_full_indiv, "BASELINE configs[4]", reps=2, warm=1,
                                     host_barrier=host_barrier, e2e_devices=e2e_devices, sampled=True)

    units_step = 2.0 * snps * indiv * ncol
    value = units_step * args.steps / (total_ms * 1e-3)

    # Tear down in an order that never leaves torch holding events on the library's stream:
    # pinned and device panels first (the host allocator records an event on every stream that
    # used a pinned block), then the process group, then the handle (which destroys the stream).
    del hBn_full, hBt, Bn, Bt, Cn, Ct
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    if hasattr(torch._C, "_host_emptyCache"):
        torch._C._host_emptyCache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    sh._stream = None
    del stream
    lib.free_compressed(C.byref(handle))
    if rank != 0:
        return

    hbm, _bf16, peak_src = peaks()
    bytes_n = dgemm_bytes(indiv, s_loc, ncol)
    bytes_t = dgemm_bytes(s_loc, indiv, ncol)
    # dominant kernel = the main dgemm_compressed kernel ('n' and 't' launches of the same family)
    k_all_ms = (sum(k_n) + sum(k_t))
    k_launches = max(1, len(k_n) + len(k_t))
    avg_bytes = (bytes_n * len(k_n) + bytes_t * len(k_t)) / k_launches
    achieved = avg_bytes / (k_all_ms / k_launches * 1e-3) / 1e9 if k_all_ms > 0 else 0.0
    roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s",
                "frac": achieved / hbm, "traffic": None, "peak_source": peak_src,
                "kernel": "dgemm_compressed main kernel (avg of 'n' and 't' launches)",
                "n_ms": statistics.mean(k_n) if k_n else None,
                "t_ms": statistics.mean(k_t) if k_t else None,
                "n_gbs": bytes_n / (statistics.mean(k_n) * 1e-3) / 1e9 if k_n else None,
                "t_gbs": bytes_t / (statistics.mean(k_t) * 1e-3) / 1e9 if k_t else None,
                "share_of_step": k_all_ms / max(1, args.steps) / (total_ms / args.steps)
                if got == 2 * args.steps else None}
    tfile = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tfile):
        try:
            roofline["traffic"] = json.load(open(tfile)).get("dgemm_bytes_per_launch")
        except Exception:
            pass

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args), "snps": snps, "indiv": indiv, "ncol": ncol,
                       "step": "one 'n' call + one 't' call (+ all-reduce of indiv x ncol for 'n' when N>1)",
                       "l2": "inputs larger than L2 (packed shard per direction = %.1f GB)"
                             % (s_loc * bs / 1e9),
                       "dgemm_mode": int(os.environ.get("MB200_DGEMM_MODE", "-1")),
                       "setup_s": setup_s},
            "roofline": roofline,
            "e2e": e2e,
            "gpu_launches": int(launches), "clocks": clocks}

    if small is not None:
        line["small"] = small
    if wide is not None:
        line["wide_b"] = wide
    if pcg is not None:
        line["pcg"] = pcg
    if not args.no_grm and world == 1:
        line["grm"] = run_grm(args, lib, api, dev)
    if grm_part is not None:
        line["grm"] = grm_part
    if grm_full is not None:
        line["grm_full"] = grm_full
    if not args.no_cpu:
        try:
            line["cpu_baseline"] = run_cpu_baseline(args)
        except Exception as exc:  # the CPU checker must never take the GPU line down
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                                    "sample": f"failed: {exc}"}
        if isinstance(line.get("small"), dict):
            try:
                line["small"]["cpu_baseline"] = run_small_cpu_baseline()
            except Exception as exc:
                line["small"]["cpu_baseline"] = {"value": None, "kind": "reference", "sample": f"failed: {exc}"}
        if rank == 0 and world == 1 and isinstance(line.get("grm"), dict):
            try:
                line["grm"]["cpu_baseline"] = run_grm_cpu_baseline()
            except Exception as exc:
                line["grm"]["cpu_baseline"] = {"value": None, "unit": "TOPS", "cores": 0,
                                               "kind": "reference", "sample": f"failed: {exc}"}
    print(json.dumps(line))



def _wall_steps(fn, steps, warm):
    """Wall-clock seconds per call of a BLOCKING host call (the C-ABI entry points return after the
    result is in host memory), mean and best over `steps` calls after `warm` warm-ups."""
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(steps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return sum(ts) / len(ts), min(ts)


def host_link_probe(devices, mb=64, reps=4):
    """Aggregate pinned host <-> device copy bandwidth with ALL `devices` copying at once (one
    stream per device, 64 MB buffers): the floor of any e2e number is bytes / this.  On cloud
    boxes the GPUs often share host links, so the aggregate does not grow with the device count."""
    import torch
    devs = sorted(set(devices))
    out = {}
    try:
        hb = [torch.empty(mb << 20, dtype=torch.uint8).pin_memory() for _ in devs]
        db = [torch.empty(mb << 20, dtype=torch.uint8, device=f"cuda:{d}") for d in devs]
        st = [torch.cuda.Stream(device=f"cuda:{d}") for d in devs]
        for name in ("h2d", "d2h"):
            best = 0.0
            for it in range(reps + 1):
                for d in devs:
                    torch.cuda.synchronize(d)
                t0 = time.perf_counter()
                for i, d in enumerate(devs):
                    with torch.cuda.stream(st[i]):
                        if name == "h2d":
                            db[i].copy_(hb[i], non_blocking=True)
                        else:
                            hb[i].copy_(db[i], non_blocking=True)
                for d in devs:
                    torch.cuda.synchronize(d)
                dt = time.perf_counter() - t0
                if it:
                    best = max(best, len(devs) * (mb << 20) / dt / 1e9)
            out[name + "_gbs_aggregate"] = best
        out["devices"] = devs
        out["how"] = f"{mb} MB pinned buffer per device, all devices at once, best of {reps}, wall clock"
    except Exception as exc:
        out["error"] = repr(exc)
    return out


def run_e2e(args, lib, api, handle, devices, seed, hBn, hBt, Cn_dev, Ct_dev, steps):
    """dgemm_compressed() 'n' + 't' through the C ABI with host panels, one process.  handle: an
    existing single-device handle, or None = create a handle of the same simulated matrix over
    `devices` (the MIRACULIX_NGPUS path).  Pinned panels (headline), pageable panels beside them,
    and a parity check of both against the device-resident arm's result."""
    import numpy as np
    import torch
    snps, indiv, ncol = args.snps, args.indiv, args.ncol
    own = handle is None
    if own:
        api.set_devices(devices)
        try:
            handle = api.simulate_compressed(snps, indiv, seed, ncol)
        finally:
            api.set_devices(None)
    nsh = lib.mb200_shard_count(handle)
    hCn = torch.empty((ncol, indiv), dtype=torch.float64).pin_memory()
    hCt = torch.empty((ncol, snps), dtype=torch.float64).pin_memory()

    def call(tr, B, k, Cout, m):
        lib.dgemm_compressed(tr, handle, ncol, B, k, Cout, m)
        if lib.mb200_last_status() != 0:
            raise SystemExit("dgemm_compressed failed: " + lib.mb200_last_message().decode())

    def step_pinned():
        call(b"N", hBn.data_ptr(), snps, hCn.data_ptr(), indiv)
        call(b"T", hBt.data_ptr(), indiv, hCt.data_ptr(), snps)

    launches0 = api.kernel_launches()
    mean_s, best_s = _wall_steps(step_pinned, steps, 2)
    launches = (api.kernel_launches() - launches0) // (steps + 2)
    n_s, _ = _wall_steps(lambda: call(b"N", hBn.data_ptr(), snps, hCn.data_ptr(), indiv), steps, 0)
    t_s, _ = _wall_steps(lambda: call(b"T", hBt.data_ptr(), indiv, hCt.data_ptr(), snps), steps, 0)

    # parity: the e2e result against the device-resident arm (same matrix, same B)
    def rel(a, b):
        return float((a - b).abs().max() / b.abs().max())
    par_n = rel(hCn, Cn_dev)
    par_t = rel(hCt[:, :Ct_dev.shape[1]], Ct_dev)       # rank 0's SNP rows come first

    # pageable panels: plain numpy arrays, as a Julia Matrix{Float64} / Fortran array would be
    pBn, pBt = hBn.numpy().copy(), hBt.numpy().copy()
    pCn, pCt = np.empty((ncol, indiv)), np.empty((ncol, snps))

    def step_pageable():
        call(b"N", pBn.ctypes.data, snps, pCn.ctypes.data, indiv)
        call(b"T", pBt.ctypes.data, indiv, pCt.ctypes.data, snps)
    pmean_s, pbest_s = _wall_steps(step_pageable, steps, 2)
    par_p = max(rel(torch.from_numpy(pCn), hCn), rel(torch.from_numpy(pCt), hCt))
    if own:
        lib.free_compressed(C.byref(handle))
    link = host_link_probe(devices)
    units = 2.0 * snps * indiv * ncol
    nbytes = 8 * ncol * (snps + indiv)
    floor_ms = None
    if link.get("h2d_gbs_aggregate") and link.get("d2h_gbs_aggregate"):
        # 'n' moves snps x ncol up and indiv x ncol down, 't' the reverse; the calls are sequential
        floor_ms = (8e-6 * ncol * (snps / link["h2d_gbs_aggregate"] + indiv / link["d2h_gbs_aggregate"]) +
                    8e-6 * ncol * (indiv / link["h2d_gbs_aggregate"] + snps / link["d2h_gbs_aggregate"]))
    return {"value": units / mean_s, "unit": UNIT, "h2d_bytes_per_step": nbytes,
            "d2h_bytes_per_step": nbytes, "steps": steps, "ms_per_step": mean_s * 1e3,
            "best_ms_per_step": best_s * 1e3, "n_ms": n_s * 1e3, "t_ms": t_s * 1e3,
            "timing": "wall clock around the blocking calls, mean of %d steps after 2 warm-ups" % steps,
            "api": ("dgemm_compressed() C-ABI, pinned host B/C" if nsh == 1 else
                    "dgemm_compressed() C-ABI on ONE sharded handle: one process, %d device shards %s, "
                    "peer-memory all-gather ('t') / reduce-scatter ('n'), pinned host B/C" % (nsh, devices)),
            "devices": devices, "shards": nsh, "gpu_launches_per_step": int(launches),
            "host_link": link, "host_link_floor_ms_per_step": floor_ms,
            "pageable": {"value": units / pmean_s, "ms_per_step": pmean_s * 1e3,
                         "best_ms_per_step": pbest_s * 1e3, "vs_pinned": pmean_s / mean_s,
                         "what": "the same calls with pageable numpy panels (staged through the "
                                 "library's pinned ring)"},
            "parity": {"n_vs_device_arm": par_n, "t_vs_device_arm": par_t, "pageable_vs_pinned": par_p,
                       "ok": bool(par_n <= 1e-10 and par_t <= 1e-10 and par_p <= 1e-10),
                       "bar": "max|dC|/max|C| <= 1e-10"}}


def run_wide(args, lib, api, dev):
    """dgemm_compressed with B wider than one MMA's digit rows (examples/gblup/calculate_gblup.jl:89-93
    calls with n + p = 50 columns): device-resident 'n' + 't' on the configs[2] matrix.  The column
    groups run in one launch and share the packed tiles in L2; DRAM should see the matrix once."""
    import torch
    snps, indiv, ncol = args.snps, args.indiv, args.wide_ncol
    h = api.simulate_compressed(snps, indiv, 1000, ncol)
    g = torch.Generator(device=dev)
    g.manual_seed(3)
    Bn = torch.randn((ncol, snps), dtype=torch.float64, device=dev, generator=g)
    Bt = torch.randn((ncol, indiv), dtype=torch.float64, device=dev, generator=g)
    Cn = torch.empty((ncol, indiv), dtype=torch.float64, device=dev)
    Ct = torch.empty((ncol, snps), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()

    def step():
        api._check(lib.mb200_dgemm_compressed_device(0, h, ncol, C.c_void_p(Bn.data_ptr()), snps,
                                                     C.c_void_p(Cn.data_ptr()), indiv), "dgemm n")
        api._check(lib.mb200_dgemm_compressed_device(1, h, ncol, C.c_void_p(Bt.data_ptr()), indiv,
                                                     C.c_void_p(Ct.data_ptr()), snps), "dgemm t")
    reps = 5
    for _ in range(2):
        step()
    lib.mb200_sync(h)
    ms_arr, tr_arr = (C.c_float * (2 * reps))(), (C.c_int * (2 * reps))()
    t0 = time.perf_counter()
    for _ in range(reps):
        step()
    lib.mb200_sync(h)
    sec = (time.perf_counter() - t0) / reps
    got = lib.mb200_dgemm_kernel_times(h, 2 * reps, ms_arr, tr_arr)
    k_ms = sum(ms_arr[i] for i in range(got)) / max(1, got)
    sl = C.c_int(0)
    lib.mb200_dgemm_last_precision(h, C.byref(sl))
    lib.free_compressed(C.byref(h))
    del Bn, Bt, Cn, Ct
    torch.cuda.empty_cache()
    hbm, _b, _s = peaks()
    alg = (dgemm_bytes(indiv, snps, ncol) + dgemm_bytes(snps, indiv, ncol)) / 2
    groups = -(-ncol // (80 // max(1, sl.value)))
    return {"workload": f"dgemm_compressed 'n'+'t', ncol={ncol}, {snps} SNPs x {indiv} individuals, device-resident",
            "ms_per_step": sec * 1e3, "main_kernel_ms": k_ms, "digits": sl.value, "column_groups": groups,
            "value": 2.0 * snps * indiv * ncol / sec, "unit": UNIT,
            "roofline": {"bound": "hbm", "achieved": alg / (k_ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                         "frac": alg / (k_ms * 1e-3) / 1e9 / hbm,
                         "note": "algorithmic bytes = the packed matrix ONCE + B + C; the launch does "
                                 f"{groups}x the tensor work of a {80 // max(1, sl.value)}-column call"}}


def run_small(lib, api):
    """BASELINE configs[0]: dgemm_compressed 'n', ncol = 1, 10k SNPs x 1k individuals -- the
    reference's own test case (tests/dgemm_compressed); latency of the C-ABI call with pageable
    host vectors.  Reported, not a roofline case (the matrix is 2.5 MB).  The default configuration
    routes a call this small to the one-kernel path (csrc/dgemm_small.cu); the same call through
    the tcgen05 digit-slice path (small limit 0) is timed beside it and the two results compared."""
    import numpy as np
    snps, indiv = 10_000, 1_000
    h = api.simulate_compressed(snps, indiv, 5, 1)
    rng = np.random.default_rng(1)
    B = rng.standard_normal(snps)
    res, outs, ok = {}, {}, True
    for name, limit in (("default", 4.0e7), ("tensor_core", 0.0)):
        lib.mb200_set_dgemm_small_limit(limit)
        Cout = np.empty(indiv)
        ts = []
        for it in range(120):
            t0 = time.perf_counter()
            lib.dgemm_compressed(b"N", h, 1, B.ctypes.data, snps, Cout.ctypes.data, indiv)
            if it >= 20:
                ts.append(time.perf_counter() - t0)
        ok = ok and lib.mb200_last_status() == 0
        res[name], outs[name] = ts, Cout
    lib.mb200_set_dgemm_small_limit(4.0e7)
    lib.free_compressed(C.byref(h))
    ts = res["default"]
    diff = float(np.abs(outs["default"] - outs["tensor_core"]).max() / np.abs(outs["tensor_core"]).max())
    return {"workload": "dgemm_compressed 'n', ncol=1, 10000 SNPs x 1000 individuals (BASELINE configs[0])",
            "ms_median": statistics.median(ts) * 1e3, "ms_best": min(ts) * 1e3, "calls": len(ts),
            "value": snps * indiv / statistics.median(ts), "unit": UNIT, "ok": ok and diff <= 1e-10,
            "path": "one fused fp64 kernel (csrc/dgemm_small.cu), the default below 4e7 genotype-column products",
            "tensor_core_path": {"ms_median": statistics.median(res["tensor_core"]) * 1e3,
                                 "ms_best": min(res["tensor_core"]) * 1e3,
                                 "max_rel_diff_of_results": diff},
            "api": "dgemm_compressed() C-ABI, pageable host vectors, wall clock per call"}


def run_pcg(args, lib, api, dev, rank, world, host_barrier=lambda: None, e2e_devices=None):
    """BASELINE configs[3]: conjugate gradients on (Zc Zc^T + lambda I) x = b, 500k SNPs x 200k
    individuals, packed matrix resident and SNP-sharded over the ranks; one iteration = one 't' and
    one 'n' product with a single column + ONE all-reduce of indiv doubles (mb200_cg_solve_device).
    Every rank runs it (collective); rank 0 reports."""
    import torch
    import torch.distributed as dist
    from miraculix_b200.sharded import ShardedCompressed
    snps, indiv, iters = args.pcg_snps, args.pcg_indiv, args.pcg_iters
    sh = ShardedCompressed(None, snps, indiv, rank, world)
    s_loc, bs = sh.s_count, (indiv + 3) // 4
    bed = torch.empty(s_loc * bs, dtype=torch.uint8, device=dev)
    api._check(lib.mb200_simulate_plink_rows_device(C.c_void_p(bed.data_ptr()), s_loc, indiv, 3000,
                                                    sh.s_begin), "simulate")
    h = C.c_void_p(None)
    api._check(lib.mb200_plink2compressed_device(C.c_void_p(bed.data_ptr()), s_loc, indiv, None, 1,
                                                 C.byref(h)), "plink2compressed_device")
    del bed
    torch.cuda.empty_cache()
    sh.handle = h
    stream = sh.stream()
    import numpy as np
    hb = np.random.default_rng(11).standard_normal(indiv)
    b = torch.from_numpy(hb).to(dev)
    x = torch.zeros(indiv, dtype=torch.float64, device=dev)
    lam = 0.05 * snps
    sh.cg_solve(b, x, lam=lam, tol=0.0, max_iter=2)  # warm-up (workspace, NCCL channel)
    x.zero_()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = api.kernel_launches()
    e0.record(stream)
    it, res = sh.cg_solve(b, x, lam=lam, tol=0.0, max_iter=iters)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    launches = api.kernel_launches() - launches0
    bnorm = float(b.norm().item())
    x_dev = x.cpu().numpy()
    # e2e: mb200_cg_solve with HOST vectors, one process; at N > 1 rank 0 alone on a sharded handle
    torch.cuda.synchronize()
    host_barrier()
    e2e = None
    if rank == 0:
        devices = e2e_devices or list(range(world))
        own = len(devices) > 1
        hh = h
        if own:
            api.set_devices(devices)
            try:
                hh = api.simulate_compressed(snps, indiv, 3000, 1)
            finally:
                api.set_devices(None)
        api.cg_solve(hh, hb, np.zeros(indiv), lam=lam, tol=0.0, max_iter=2)
        t0 = time.perf_counter()
        xe, ite, rese = api.cg_solve(hh, hb, np.zeros(indiv), lam=lam, tol=0.0, max_iter=iters)
        sec = time.perf_counter() - t0
        if own:
            lib.free_compressed(C.byref(hh))
        e2e = {"ms_per_iteration": sec * 1e3 / (ite + 1), "iterations": ite,
               "value": 2.0 * snps * indiv / (sec / (ite + 1)), "unit": UNIT, "devices": devices,
               "api": "mb200_cg_solve() C-ABI, host vectors, one process"
                      + (", sharded handle over %d devices" % len(devices) if own else ""),
               "x_vs_device_arm": float(np.abs(xe - x_dev).max() / max(np.abs(x_dev).max(), 1e-300)),
               "residual_reduction": rese / bnorm}
    host_barrier()
    sh._stream = None
    del stream
    lib.free_compressed(C.byref(h))
    if rank != 0:
        return None
    hbm, _bf16, _src = peaks()
    products = it + 1                                     # the initial G x0 and one per iteration
    ms_iter = float(ms.item()) / products
    alg_bytes = 2.0 * snps * bs + 8.0 * 2 * (snps + indiv)  # both orientations once + vectors
    gbs = alg_bytes / world / (ms_iter * 1e-3) / 1e9
    return {"workload": f"CG on (ZcZc^T + {lam:g} I) x = b, {snps} SNPs x {indiv} individuals "
                        f"(BASELINE configs[3]), packed matrix resident, SNP-sharded over {world} GPU(s)",
            "iterations": it, "ms_per_iteration": ms_iter,
            "value": 2.0 * snps * indiv / (ms_iter * 1e-3), "unit": UNIT,
            "residual_reduction": res / bnorm, "gpu_launches": int(launches), "e2e": e2e,
            "roofline": {"bound": "hbm", "achieved": gbs, "peak": hbm, "unit": "GB/s per GPU",
                         "frac": gbs / hbm,
                         "note": "algorithmic bytes = both packed orientations once per iteration "
                                 "(ncol = 1: 't' then 'n'), per GPU"}}



def int8_peak_file():
    """(peak TOP/s, source).  profiles/int8_peak.json holds the cuBLASLt int8 rate MEASURED on this
    pool (written from an earlier bench run); the nominal dense figure is the fallback."""
    f = os.path.join(ROOT, "profiles", "int8_peak.json")
    try:
        d = json.load(open(f))
        return float(d["nominal_dense_tops"]), float(d["measured_cublaslt_tops"]), "profiles/int8_peak.json"
    except Exception:
        return 4500.0, None, "nominal int8 dense (no profiles/int8_peak.json)"


def _mem_available_bytes():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable:"):
                return int(ln.split()[1]) * 1024
    except Exception:
        pass
    return 0


def grm_row_sums_exact(lib, api, h, snps, indiv):
    """m = Z (Z^T 1) and its total, exactly, through dgemm_compressed itself (not centred, 8 digits:
    integer-exact), for the sampled parity check of the GRM legs."""
    import numpy as np
    api.set_options(use_gpu=True, not_center=True)
    try:
        colsum = api.dgemm_compressed("T", h, np.ones((indiv, 1), order="F"), snps, indiv)
        m = api.dgemm_compressed("N", h, colsum, snps, indiv).ravel()
    finally:
        api.set_options(use_gpu=True, not_center=False)
    return m, float(m.sum())


def grm_sampled_parity(G, hx, indiv, m, total, c, samples=32, seed=3):
    """Sampled entries (both triangles) of a centred+scaled GRM against direct int64 dot products of
    the decoded packed rows and the update of crossproduct.jl:91-106 in numpy long double."""
    import numpy as np
    rng = np.random.default_rng(seed)
    n = indiv
    ii = rng.integers(0, n, samples)
    jj = rng.integers(0, n, samples)
    ii[0], jj[0] = 0, n - 1
    ii[1], jj[1] = n - 1, n - 1
    ii[2] = jj[2]                                  # a diagonal entry
    shifts = np.arange(4, dtype=np.uint8) * 2

    def row(i):
        b = np.asarray(hx[i])
        return ((b[:, None] >> shifts[None, :]) & 3).reshape(-1).astype(np.int64)
    worst, gmax = 0.0, float(np.abs(G[ii, jj]).max())
    for i, j in zip(ii.tolist(), jj.tolist()):
        cnt = int(row(i) @ row(j))
        want = (np.longdouble(cnt) - (np.longdouble(m[i]) + np.longdouble(m[j])) / n
                + np.longdouble(total) / n / n) / np.longdouble(c)
        for got in (G[i, j], G[j, i]):
            worst = max(worst, abs(float(got - want)))
    gmax = max(gmax, 1e-300)
    return {"samples": int(samples), "max_abs_err": worst, "max_abs_value": gmax,
            "rel": worst / gmax, "ok": bool(worst <= 1e-10 * gmax),
            "bar": "|dG| <= 1e-10 max|G| on sampled (i,j) and (j,i) against int64 dot products"}


def grm_e2e(lib, api, hx, kcols, indiv, hf, devices, pinned, reps, parity=None):
    """mb200_grm() host call (= snp_multiply_gpu + the fused crossproduct.jl:83-110 epilogue) over
    `devices`, one process: packed matrix and fp64 result in HOST memory, both PCIe legs and all
    device allocations inside the timed region (wall clock; the call blocks)."""
    import numpy as np
    import torch
    need = 8 * indiv * indiv
    if need * 1.15 > _mem_available_bytes():
        return {"skipped": "host memory: result needs %.1f GB, %.1f GB available"
                           % (need / 1e9, _mem_available_bytes() / 1e9)}, None
    if pinned:
        hans_t = torch.empty(indiv * indiv, dtype=torch.float64).pin_memory()
        hans = hans_t.numpy().reshape(indiv, indiv)
        hx_t = torch.from_numpy(hx).pin_memory()          # the packed input is page-locked too
        hx = hx_t.numpy()
    else:
        hans = np.empty((indiv, indiv))
        hans.fill(0.0)                                   # first touch outside the timed region
    api.set_devices(devices if len(devices) > 1 else None)
    ts = []
    try:
        for it in range(reps + 1):
            t0 = time.perf_counter()
            api._check(lib.mb200_grm(C.c_void_p(hx.ctypes.data), kcols, indiv, C.c_void_p(hans.ctypes.data),
                                     False, C.c_void_p(hf.ctypes.data), True), "mb200_grm")
            if it >= 1 or reps == 0:
                ts.append(time.perf_counter() - t0)
    finally:
        api.set_devices(None)
    t = api.crossprod_last_timing()
    out = {"ms": statistics.mean(ts) * 1e3, "best_ms": min(ts) * 1e3, "device_pipeline_ms": t["gemm_ms"],
           "h2d_bytes_per_step": int(hx.nbytes + 8 * kcols), "d2h_bytes_per_step": int(need),
           "devices": devices, "host_memory": "pinned" if pinned else "pageable",
           "api": "mb200_grm() C-ABI host call, one process, %d device(s), %d timed call(s) after 1 warm-up"
                  % (len(devices), len(ts))}
    if parity is not None:
        m, total, c = parity
        out["sampled_parity"] = grm_sampled_parity(hans, hx, indiv, m, total, c)
    out["symmetric"] = bool(np.array_equal(hans[:257, :4096], hans[:4096, :257].T))
    return out, hans


def run_grm_partition(args, lib, api, dev, rank, world, snps, indiv, label, reps=None, warm=None,
                      host_barrier=lambda: None, e2e_devices=None, sampled=False):
    """configs[1]/[4] over N GPUs: 2-D output-tile partition, NO communication -- every rank holds
    the packed matrix and computes the 128-row tiles t with t % N == rank of the centred GRM
    (mb200_crossprod_partition_device).  Time = max over ranks of the device time."""
    import torch
    import torch.distributed as dist
    warm = args.warmup if warm is None else warm
    reps = warm + (max(3, min(args.steps, 10)) if reps is None else reps)
    bs = (indiv + 3) // 4
    torch.cuda.empty_cache()
    bed = torch.empty(snps * bs, dtype=torch.uint8, device=dev)
    api._check(lib.mb200_simulate_plink_device(C.c_void_p(bed.data_ptr()), snps, indiv, 4242), "simulate")
    h = C.c_void_p(None)
    api._check(lib.mb200_plink2compressed_device(C.c_void_p(bed.data_ptr()), snps, indiv, None, 1,
                                                 C.byref(h)), "plink2compressed_device")
    del bed
    torch.cuda.empty_cache()  # the library allocates from the driver's pool, not torch's cache
    lay = api.get_layout(h)
    kcols = lay.pitch_indiv_major * 4
    f = torch.zeros(kcols, dtype=torch.float64, device=dev)
    f[:snps].copy_(torch.frombuffer(bytearray(api.get_compressed_freq(h, snps).tobytes()),
                                    dtype=torch.float64))
    lrows = lib.mb200_crossprod_partition_rows(indiv, rank, world)
    panel = torch.empty(max(1, lrows) * indiv, dtype=torch.float64, device=dev)
    gemm, epi, ops = [], [], 0.0
    for it in range(reps):
        dist.barrier()
        api._check(lib.mb200_crossprod_partition_device(C.c_void_p(lay.d_indiv_major), kcols, indiv, 0, 1,
                                                        C.c_void_p(f.data_ptr()), 1, rank, world,
                                                        C.c_void_p(panel.data_ptr())), "crossprod_partition")
        t = api.crossprod_last_timing()
        if it >= warm:
            gemm.append(t["gemm_ms"])
            epi.append(t["epilogue_ms"])
            ops = t["executed_ops"]
    torch.cuda.synchronize()
    # e2e leg (rank 0, one process over all GPUs): the packed matrix goes to HOST memory first
    e2e_in = None
    if rank == 0:
        import numpy as np
        nbytes = indiv * lay.pitch_indiv_major
        hx = np.empty((indiv, lay.pitch_indiv_major), dtype=np.uint8)
        step = max(1, (1 << 30) // lay.pitch_indiv_major)
        for r0 in range(0, indiv, step):                      # 1 GiB pieces through torch's D2H
            r1 = min(indiv, r0 + step)
            hx[r0:r1] = _device_bytes(lay.d_indiv_major + r0 * lay.pitch_indiv_major,
                                      (r1 - r0) * lay.pitch_indiv_major, dev).cpu().numpy().reshape(r1 - r0, -1)
        hf = f.cpu().numpy()
        par = None
        if sampled:
            m_exact, total = grm_row_sums_exact(lib, api, h, snps, indiv)
            par = (m_exact, total, float(2.0 * (hf * (1.0 - hf)).sum()))
        e2e_in = (hx, hf, par)
    lib.free_compressed(C.byref(h))
    del panel, f
    torch.cuda.empty_cache()
    v = torch.tensor([statistics.mean(gemm), statistics.mean(epi), ops], dtype=torch.float64, device=dev)
    vmax = v.clone()
    dist.all_reduce(vmax, op=dist.ReduceOp.MAX)
    dist.all_reduce(v, op=dist.ReduceOp.SUM)
    torch.cuda.synchronize()
    host_barrier()
    e2e = None
    if rank == 0:
        hx, hf, par = e2e_in
        devices = e2e_devices or list(range(world))
        big = indiv * indiv * 8 > 8e9
        try:
            e2e, _hans = grm_e2e(lib, api, hx, kcols, indiv, hf, devices, pinned=not big,
                                 reps=1 if big else 3, parity=par)
            del _hans
        except Exception as exc:  # the leg is reported, never fatal for the headline
            e2e = {"error": repr(exc)}
        if isinstance(e2e, dict) and "ms" in e2e:
            e2e["value"] = 2.0 * indiv * indiv * snps / (e2e["ms"] * 1e-3) / 1e12
            e2e["unit"] = "TOPS"
        del hx
    host_barrier()
    if rank != 0:
        return None
    gm, ep, ops_sum = float(vmax[0]), float(vmax[1]), float(v[2])
    tot = gm + ep
    alg_ops = 2.0 * indiv * indiv * snps
    int8_peak_v, int8_meas, int8_src = int8_peak_file()
    return {"e2e": e2e, "workload": f"centred+scaled GRM, {snps} SNPs x {indiv} individuals ({label}), "
                        f"2-D output-tile partition over {world} GPUs, no communication",
            "metric": "GRM crossproduct int8 TOPS", "value": alg_ops / (tot * 1e-3) / 1e12,
            "unit": "TOPS (2*indiv^2*snps / t, full-matrix count)", "ms": tot,
            "expand_gemm_ms": gm, "epilogue_ms": ep,
            "roofline": {"bound": "tensor", "achieved": ops_sum / (gm * 1e-3) / 1e12 / world,
                         "peak": int8_peak_v, "unit": "TOP/s per GPU",
                         "frac": ops_sum / (gm * 1e-3) / 1e12 / world / int8_peak_v, "traffic": None,
                         "peak_source": int8_src, "peak_measured_library": int8_meas,
                         "note": "achieved = EXECUTED ops of all ranks / (max expand+GEMM time x N)"}}


def library_int8_peak(dev, indiv, snps):
    """cuBLASLt int8 GEMM (torch._int_mm) on this box, as the MEASURED int8 denominator SURVEY 8(d)
    asks for beside the nominal 4.5 POPS: a square 8192^3 product and the GRM's own shape as a plain
    full-matrix int8 GEMM (no expansion from 2 bits, no symmetry, no fp64 epilogue).  Library calls,
    reported only; never on the product path."""
    import torch
    out = {}
    try:
        def timed(m, k, n, reps=10):
            a = torch.randint(-2, 3, (m, k), dtype=torch.int8, device=dev)
            b = torch.randint(-2, 3, (n, k), dtype=torch.int8, device=dev).t()
            for _ in range(3):
                torch._int_mm(a, b)
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
            torch.cuda.synchronize()
            ev[0].record()
            for i in range(reps):
                torch._int_mm(a, b)
                ev[i + 1].record()
            torch.cuda.synchronize()
            ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(reps)]
            return 2.0 * m * k * n / 1e12, min(ms), statistics.mean(ms)
        tera, best, mean = timed(8192, 8192, 8192)
        out["square_8192"] = {"best_tops": tera / (best * 1e-3), "mean_tops": tera / (mean * 1e-3)}
        m = (indiv + 7) // 8 * 8
        k = (snps + 7) // 8 * 8
        tera, best, mean = timed(m, k, m)
        out["grm_shape_full_matrix"] = {"m": m, "k": k, "best_ms": best, "mean_ms": mean,
                                        "best_tops": tera / (best * 1e-3), "mean_tops": tera / (mean * 1e-3)}
        out["how"] = "torch._int_mm (cuBLASLt int8 -> int32), 10 back-to-back calls after 3 warm-ups, CUDA events"
    except Exception as e:  # reported figure only
        out["error"] = repr(e)
    return out


def _device_bytes(ptr, nbytes, dev):
    """uint8 torch view of a library-owned device buffer (no copy)."""
    import torch

    class _Holder:
        __cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (int(ptr), False),
                                    "version": 3, "strides": None}
    return torch.as_tensor(_Holder(), device=dev)


def _grm_traffic(snps, indiv):
    """DRAM bytes per crossprod_tc_kernel launch from the committed ncu capture (default shape only)."""
    tfile = os.path.join(ROOT, "profiles", "traffic.json")
    if (snps, indiv) != (50_000, 10_000) or not os.path.exists(tfile):
        return None
    try:
        return json.load(open(tfile)).get("grm_bytes_per_launch")
    except Exception:
        return None


def run_grm(args, lib, api, dev):
    """Second half of the metric: centred GRM of configs[1] (50k SNPs x 10k individuals)."""
    import torch
    snps, indiv = args.grm_snps, args.grm_indiv
    bs = (indiv + 3) // 4
    bed = torch.empty(snps * bs, dtype=torch.uint8, device=dev)
    api._check(lib.mb200_simulate_plink_device(C.c_void_p(bed.data_ptr()), snps, indiv, 4242), "simulate")
    h = C.c_void_p(None)
    api._check(lib.mb200_plink2compressed_device(C.c_void_p(bed.data_ptr()), snps, indiv, None, 1,
                                                 C.byref(h)), "plink2compressed_device")
    lay = api.get_layout(h)
    ans = torch.empty(indiv * indiv, dtype=torch.float64, device=dev)
    # individual-major value-coded rows (pitch bytes per row): K = 4 * pitch columns, pad = 0
    kcols = lay.pitch_indiv_major * 4
    f = torch.zeros(kcols, dtype=torch.float64, device=dev)
    f[:snps].copy_(torch.frombuffer(bytearray(api.get_compressed_freq(h, snps).tobytes()),
                                    dtype=torch.float64))
    gemm, epi, ops = [], [], 0.0
    for it in range(args.warmup + max(3, min(args.steps, 10))):
        api._check(lib.mb200_crossprod_device(C.c_void_p(lay.d_indiv_major), kcols, indiv, 0, 1,
                                              C.c_void_p(f.data_ptr()), 1, C.c_void_p(ans.data_ptr())),
                   "crossprod")
        t = api.crossprod_last_timing()
        if it >= args.warmup:
            gemm.append(t["gemm_ms"])
            epi.append(t["epilogue_ms"])
            ops = t["executed_ops"]
    torch.cuda.synchronize()
    # end to end through the reference-facing host call (mb200_grm = snp_multiply_gpu + the fused
    # crossproduct.jl:83-110 epilogue): packed matrix and fp64 result in HOST memory (pinned, then
    # pageable), both PCIe legs and the device allocations inside the timed region.
    e2e = None
    try:
        import numpy as np
        nbytes = indiv * lay.pitch_indiv_major
        hx = _device_bytes(lay.d_indiv_major, nbytes, dev).cpu().numpy().reshape(indiv, -1)
        hf = f.cpu().numpy()
        m_exact, total = grm_row_sums_exact(lib, api, h, snps, indiv)
        par = (m_exact, total, float(2.0 * (hf * (1.0 - hf)).sum()))
        e2e, hans = grm_e2e(lib, api, hx, kcols, indiv, hf, [0], pinned=True, reps=3, parity=par)
        e2e["value"] = 2.0 * indiv * indiv * snps / (e2e["ms"] * 1e-3) / 1e12
        e2e["unit"] = "TOPS"
        e2e["matches_device_result"] = bool(np.array_equal(hans, ans.cpu().numpy().reshape(indiv, indiv)))
        del hans
        pg, hans = grm_e2e(lib, api, hx, kcols, indiv, hf, [0], pinned=False, reps=3)
        e2e["pageable"] = {"ms": pg["ms"], "best_ms": pg["best_ms"], "vs_pinned": pg["ms"] / e2e["ms"]}
        del hx, hans
    except Exception as exc:  # the leg is reported, never fatal for the headline
        e2e = {"error": repr(exc)}
    lib.free_compressed(C.byref(h))
    alg_ops = 2.0 * indiv * indiv * snps
    gm = statistics.mean(gemm)
    tot = gm + statistics.mean(epi)
    int8_peak, _int8_meas, int8_src = int8_peak_file()
    lib_int8 = library_int8_peak(dev, indiv, snps)
    return {"library_int8": lib_int8, "e2e": e2e,
            "workload": f"centred+scaled GRM, {snps} SNPs x {indiv} individuals (BASELINE configs[1])",
            "metric": "GRM crossproduct int8 TOPS", "value": alg_ops / (tot * 1e-3) / 1e12,
            "unit": "TOPS (2*indiv^2*snps / t, full-matrix count)", "ms": tot,
            "expand_gemm_ms": gm, "epilogue_ms": statistics.mean(epi),
            "roofline": {"bound": "tensor", "achieved": ops / (gm * 1e-3) / 1e12, "peak": int8_peak,
                         "unit": "TOP/s", "frac": ops / (gm * 1e-3) / 1e12 / int8_peak,
                         "traffic": _grm_traffic(snps, indiv), "peak_source": int8_src,
                         "peak_measured_library": (lib_int8.get("square_8192") or {}).get("best_tops"),
                         "note": "achieved = EXECUTED ops (upper-triangle tiles only) / expand+GEMM time; "
                                 "peak_measured_library = cuBLASLt int8 8192^3 on this box"}}


def run_cpu_baseline(args):
    from oracle import ref
    if not ref.available():
        raise RuntimeError("oracle/_ref/ref_miraculix.so not built")
    snps, indiv = args.cpu_snps, args.cpu_indiv
    plink, f = host_sample(snps, indiv, 1)
    mean_v, best_v, cores, sec = cpu_reference_run(plink, snps, indiv, f, args.ncol, 3, 1)
    return {"value": mean_v, "best": best_v, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": f"{snps} SNPs x {indiv} individuals sub-block, ncol={args.ncol}, 'N'+'T' per step, "
                      f"3 steps after 1 warm-up, reference CPU 5codes variant 256, {cores} threads; "
                      f"{sec:.2f} s/step"}


def run_small_cpu_baseline():
    """configs[0] on the reference CPU 5codes library: 'N', ncol = 1, 10k SNPs x 1k individuals."""
    import numpy as np
    from oracle import ref, simulate
    snps, indiv = 10_000, 1_000
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)
    plink, _, f = simulate.simulate(snps, indiv, seed=5)
    ref.set_options(cores=cores, centered=True, variant=256)
    h = ref.RefCompressed(plink, snps, indiv, f, 1)
    B = np.asfortranarray(np.random.default_rng(1).standard_normal((snps, 1)))
    Cout = np.zeros((indiv, 1), order="F")
    ts = []
    for it in range(40):
        t0 = time.perf_counter()
        h.dgemm("N", B, Cout)
        if it >= 5:
            ts.append(time.perf_counter() - t0)
    h.free()
    return {"ms_median": statistics.median(ts) * 1e3, "ms_best": min(ts) * 1e3, "cores": cores,
            "value": snps * indiv / statistics.median(ts), "unit": UNIT, "kind": "reference",
            "sample": f"the whole of configs[0], 35 calls after 5 warm-ups, reference CPU 5codes variant 256, {cores} threads"}


def run_grm_cpu_baseline(snps=16_000, indiv=2_000):
    """GRM counts on the reference CPU 5codes library: dgemm_compressed('N', not
    centred) applied to B = Z^T as doubles gives Z Z^T exactly; bounded sample, all host threads."""
    import numpy as np
    from oracle import ref, simulate
    cores = os.cpu_count() or 1
    os.environ["OMP_NUM_THREADS"] = str(cores)
    plink, _, f = simulate.simulate(snps, indiv, seed=2)
    Z = simulate.plink_decode(simulate.unpack_2bit(plink, ((indiv + 3) // 4) * 4)[:, :indiv])
    B = np.asfortranarray(Z.astype(np.float64))             # snps x indiv  (= Z^T in the GRM's sense)
    ref.set_options(cores=cores, centered=False, variant=256)
    h = ref.RefCompressed(plink, snps, indiv, f, indiv)
    out = np.zeros((indiv, indiv), order="F")
    ts = []
    for it in range(3):
        t0 = time.perf_counter()
        h.dgemm("N", B, out)
        if it >= 1:
            ts.append(time.perf_counter() - t0)
    h.free()
    ref.set_options(cores=cores, centered=True, variant=256)
    exact = bool(np.array_equal(out, (B.T @ B)))
    t = sum(ts) / len(ts)
    return {"value": 2.0 * indiv * indiv * snps / t / 1e12, "unit": "TOPS", "cores": cores,
            "kind": "reference", "exact_counts": exact,
            "sample": f"{snps} SNPs x {indiv} individuals, raw counts Z Z^T via reference CPU 5codes "
                      f"dgemm_compressed('N', not centred) on B = Z^T, 2 calls after 1 warm-up, "
                      f"{cores} threads; {t:.2f} s/call"}


def main():
    args = parse_args()
    # stdout carries exactly ONE JSON line: anything libraries print while we run (NCCL's version
    # banner, torchrun notices) is routed to stderr, and the line goes to the real stdout.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import builtins
    _print = builtins.print

    def emit(*a, **kw):
        text = " ".join(str(x) for x in a) + kw.get("end", "\n")
        os.write(real_stdout, text.encode())
    builtins.print = emit
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
