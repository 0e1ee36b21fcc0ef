#!/usr/bin/env python
"""One 'n' and one 't' product with ncol = 50 and with ncol = 13 on the configs[2] matrix (for an ncu
pass that compares the DRAM bytes of the wide launch with a one-group launch)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from miraculix_b200 import api, build
snps, indiv = 1_000_000, 100_000
build.build(); lib = api.load_library(); api.set_options(use_gpu=True, not_center=False)
lib.mb200_set_dgemm_mode(1, 6)
h = api.simulate_compressed(snps, indiv, 1000, 50)
for ncol in (13, 50):
    Bn = torch.randn((ncol, snps), dtype=torch.float64, device="cuda"); Bt = torch.randn((ncol, indiv), dtype=torch.float64, device="cuda")
    Cn = torch.empty((ncol, indiv), dtype=torch.float64, device="cuda"); Ct = torch.empty((ncol, snps), dtype=torch.float64, device="cuda")
    api._check(lib.mb200_dgemm_compressed_device(0, h, ncol, Bn.data_ptr(), snps, Cn.data_ptr(), indiv), "n")
    api._check(lib.mb200_dgemm_compressed_device(1, h, ncol, Bt.data_ptr(), indiv, Ct.data_ptr(), snps), "t")
    lib.mb200_sync(h)
    ms = C.c_float(); lib.mb200_dgemm_last_kernel_ms(h, C.byref(ms)); print("ncol", ncol, "last main kernel ms", ms.value)
lib.free_compressed(C.byref(h))
