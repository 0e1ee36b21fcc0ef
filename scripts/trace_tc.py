#!/usr/bin/env python
"""MB200_TRACE=1 python scripts/trace_tc.py [snps indiv ncol] -- cycle accounting of the
tensor-core dgemm kernel (CTA 0: expander warp 0 and the MMA warp)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("MB200_TRACE", "1")
import torch
from miraculix_b200 import api, build

snps, indiv, ncol = (int(x) for x in (sys.argv[1:4] or (1_000_000, 100_000, 10)))
build.build()
lib = api.load_library()
api.set_options(use_gpu=True, not_center=False)
dev = torch.device("cuda", 0)
bs = (indiv + 3) // 4
bed = torch.empty(snps * bs, dtype=torch.uint8, device=dev)
api._check(lib.mb200_simulate_plink_device(C.c_void_p(bed.data_ptr()), snps, indiv, 1), "sim")
h = C.c_void_p(None)
api._check(lib.mb200_plink2compressed_device(C.c_void_p(bed.data_ptr()), snps, indiv, None, ncol, C.byref(h)), "init")
del bed
B = torch.randn((ncol, snps), dtype=torch.float64, device=dev)
Cn = torch.empty((ncol, indiv), dtype=torch.float64, device=dev)
names = ["z_full wait", "LDS + z_empty arrive", "a_empty wait", "expand ALU + st issue", "wait::st",
         "fence+syncwarp+arrive", "d_full wait", "epilogue"]
mnames = ["d_empty wait", "b_full wait", "a_full wait", "polls + issue + commit + votes"]
for it in range(3):
    api._check(lib.mb200_dgemm_compressed_device(0, h, ncol, C.c_void_p(B.data_ptr()), snps,
                                                 C.c_void_p(Cn.data_ptr()), indiv), "dgemm")
    ms = C.c_float()
    lib.mb200_dgemm_last_kernel_ms(h, C.byref(ms))
    out = (C.c_longlong * 32)()
    lib.mb200_debug_read_trace(out, 32)
nchunks_cta = (snps + 511) // 512 * ((indiv + 511) // 512) / 148
slots = nchunks_cta * 4  # per expander group
tot = sum(out[i] for i in range(8))
print(f"kernel {ms.value:.3f} ms; expander warp 0 total {tot} cycles; ~{slots:.0f} slots per group")
for i, n in enumerate(names):
    print(f"  {n:28s} {out[i]:12d}  {100*out[i]/tot:5.1f}%  {out[i]/slots:8.1f} cyc/slot")
mt = sum(out[16 + i] for i in range(4))
print(f"MMA warp (group 0) total {mt} cycles; {slots:.0f} slots")
for i, n in enumerate(mnames):
    print(f"  {n:28s} {out[16+i]:12d}  {100*out[16+i]/mt:5.1f}%  {out[16+i]/slots:8.1f} cyc/slot")
lib.free_compressed(C.byref(h))

