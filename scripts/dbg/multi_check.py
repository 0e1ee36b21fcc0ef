#!/usr/bin/env python
"""Sharded handle vs single handle of the same simulated matrix: where do results differ?"""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from miraculix_b200 import api, build
snps, indiv, ncol, world = (int(x) for x in (sys.argv[1:5] or (1_000_000, 100_000, 10, 8)))
build.build(); lib = api.load_library(); api.set_options(use_gpu=True, not_center=False)
ndev = torch.cuda.device_count()
h1 = api.simulate_compressed(snps, indiv, 1000, ncol)
api.set_devices([g % ndev for g in range(world)])
hm = api.simulate_compressed(snps, indiv, 1000, ncol)
api.set_devices(None)
rng = np.random.default_rng(7)
for rep in range(3):
    for tr, k, m in (("N", snps, indiv), ("T", indiv, snps)):
        B = np.asfortranarray(rng.standard_normal((k, ncol)))
        a = api.dgemm_compressed(tr, h1, B, snps, indiv)
        b = api.dgemm_compressed(tr, hm, B, snps, indiv)
        d = np.abs(a - b)
        rel = d.max() / np.abs(a).max()
        bad_rows = np.flatnonzero(d.max(axis=1) > 1e-9 * np.abs(a).max())
        sl = C.c_int(0)
        prec = [(lib.mb200_dgemm_last_precision(s, C.byref(sl)), sl.value) for s, _, _ in api.shards(hm)]
        print(f"rep {rep} {tr}: rel {rel:.3e}; bad rows {bad_rows.size}"
              + (f" in [{bad_rows.min()}, {bad_rows.max()}]" if bad_rows.size else "") + f"; shard precision {prec[:3]}", flush=True)
        if bad_rows.size:
            # which slices?
            edges = [indiv * g // world // 32 * 32 for g in range(world)] + [indiv]
            if tr == "N":
                print("   bad rows per slice:", [int(((bad_rows >= edges[g]) & (bad_rows < edges[g + 1])).sum()) for g in range(world)])
            print("   bad cols:", np.flatnonzero(d.max(axis=0) > 1e-9 * np.abs(a).max()).tolist())
api.free_compressed(hm); api.free_compressed(h1)
