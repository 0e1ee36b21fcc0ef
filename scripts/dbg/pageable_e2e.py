#!/usr/bin/env python
"""dgemm_compressed() e2e at configs[2] with pinned vs pageable host panels (single GPU)."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from miraculix_b200 import api, build
snps, indiv, ncol = (int(x) for x in (sys.argv[1:4] or (1_000_000, 100_000, 10)))
build.build(); lib = api.load_library(); api.set_options(use_gpu=True, not_center=False)
dev = torch.device("cuda", 0); bs = (indiv + 3) // 4
bed = torch.empty(snps * bs, dtype=torch.uint8, device=dev)
api._check(lib.mb200_simulate_plink_device(C.c_void_p(bed.data_ptr()), snps, indiv, 1), "sim")
h = C.c_void_p(None)
api._check(lib.mb200_plink2compressed_device(C.c_void_p(bed.data_ptr()), snps, indiv, None, ncol, C.byref(h)), "init")
del bed
def bufs(pinned):
    mk = (lambda *s: torch.randn(*s, dtype=torch.float64).pin_memory()) if pinned else (lambda *s: torch.randn(*s, dtype=torch.float64))
    return mk(ncol, snps), mk(ncol, indiv), mk(ncol, indiv), mk(ncol, snps)
for pinned in (True, False, True, False):
    Bn, Bt, Cn, Ct = bufs(pinned)
    ts = []
    for it in range(6):
        t0 = time.perf_counter()
        lib.dgemm_compressed(b"N", h, ncol, Bn.data_ptr(), snps, Cn.data_ptr(), indiv)
        t1 = time.perf_counter()
        lib.dgemm_compressed(b"T", h, ncol, Bt.data_ptr(), indiv, Ct.data_ptr(), snps)
        t2 = time.perf_counter()
        if it: ts.append((t1 - t0, t2 - t1))
    n = min(t[0] for t in ts) * 1e3; t = min(t[1] for t in ts) * 1e3
    print(f"{'pinned  ' if pinned else 'pageable'}: N {n:.2f} ms  T {t:.2f} ms  step {n + t:.2f} ms (best of 5)", flush=True)
lib.free_compressed(C.byref(h))
