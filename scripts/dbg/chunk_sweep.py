#!/usr/bin/env python
"""e2e of dgemm_compressed() on ONE shard-sized handle (the 8-GPU shard of configs[2] by default)
for different pipeline chunk counts: where does cutting the host panel start to pay?"""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from miraculix_b200 import api, build
snps, indiv, ncol = (int(x) for x in (sys.argv[1:4] or (125_000, 100_000, 10)))
build.build(); lib = api.load_library(); api.set_options(use_gpu=True, not_center=False)
h = api.simulate_compressed(snps, indiv, 1000, ncol)
Bn = torch.randn(ncol, snps, dtype=torch.float64).pin_memory(); Bt = torch.randn(ncol, indiv, dtype=torch.float64).pin_memory()
Cn = torch.empty(ncol, indiv, dtype=torch.float64).pin_memory(); Ct = torch.empty(ncol, snps, dtype=torch.float64).pin_memory()
dBn, dBt = Bn.cuda(), Bt.cuda(); dCn, dCt = Cn.cuda(), Ct.cuda()
def best(fn, reps=12):
    ts = []
    for i in range(reps + 2):
        t0 = time.perf_counter(); fn(); 
        if i >= 2: ts.append(time.perf_counter() - t0)
    return min(ts) * 1e3, sorted(ts)[len(ts) // 2] * 1e3
def dev_n():
    lib.mb200_dgemm_compressed_device(0, h, ncol, dBn.data_ptr(), snps, dCn.data_ptr(), indiv); lib.mb200_sync(h)
def dev_t():
    lib.mb200_dgemm_compressed_device(1, h, ncol, dBt.data_ptr(), indiv, dCt.data_ptr(), snps); lib.mb200_sync(h)
print("device-resident: N %.3f ms  T %.3f ms (best)" % (best(dev_n)[0], best(dev_t)[0]))
for chunks in (1, 2, 3, 4, 6):
    lib.mb200_set_pipeline(chunks, 0.0)
    n = best(lambda: lib.dgemm_compressed(b"N", h, ncol, Bn.data_ptr(), snps, Cn.data_ptr(), indiv))
    t = best(lambda: lib.dgemm_compressed(b"T", h, ncol, Bt.data_ptr(), indiv, Ct.data_ptr(), snps))
    print("chunks %d: N best %.3f median %.3f | T best %.3f median %.3f | step %.3f" % (chunks, n[0], n[1], t[0], t[1], n[0] + t[0]), flush=True)
lib.mb200_set_pipeline(4, 1.0e7)   # default policy: chunk count from the measured link rate
for rep in range(2):
    n = best(lambda: lib.dgemm_compressed(b"N", h, ncol, Bn.data_ptr(), snps, Cn.data_ptr(), indiv))
    t = best(lambda: lib.dgemm_compressed(b"T", h, ncol, Bt.data_ptr(), indiv, Ct.data_ptr(), snps))
    print("default policy: N best %.3f median %.3f | T best %.3f median %.3f | step %.3f" % (n[0], n[1], t[0], t[1], n[0] + t[0]), flush=True)
lib.free_compressed(C.byref(h))
