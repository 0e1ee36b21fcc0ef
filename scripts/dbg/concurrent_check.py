#!/usr/bin/env python
"""Ordinary single-device handles on ONE GPU driven from several host threads at the same time:
do concurrent dgemm_tc kernels disturb each other?  usage: snps indiv ncol nthreads [dirs=NT] [reps]"""
import ctypes as C, os, sys, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from miraculix_b200 import api, build
snps, indiv, ncol, nthreads = (int(x) for x in (sys.argv[1:5] or (125_000, 100_000, 10, 4)))
dirs = (sys.argv[5:6] or ["NT"])[0]
reps = int((sys.argv[6:7] or ["6"])[0])
build.build(); lib = api.load_library(); api.set_options(use_gpu=True, not_center=False)
hs = [api.simulate_compressed(snps, indiv, 1000 + i, ncol) for i in range(nthreads)]
rng = np.random.default_rng(7)
B = {"N": np.asfortranarray(rng.standard_normal((snps, ncol))), "T": np.asfortranarray(rng.standard_normal((indiv, ncol)))}
ref = [{d: api.dgemm_compressed(d, h, B[d], snps, indiv) for d in dirs} for h in hs]
nbad = 0
for rep in range(reps):
    out = [None] * nthreads
    def work(i):
        out[i] = {d: api.dgemm_compressed(d, hs[i], B[d], snps, indiv) for d in dirs}
    th = [threading.Thread(target=work, args=(i,)) for i in range(nthreads)]
    [t.start() for t in th]; [t.join() for t in th]
    for i in range(nthreads):
        for d in dirs:
            diff = np.abs(out[i][d] - ref[i][d]); scale = np.abs(ref[i][d]).max()
            if diff.max() > 1e-10 * scale:
                nbad += 1
                rows = np.flatnonzero(diff.max(axis=1) > 1e-12 * scale)
                runs = np.split(rows, np.flatnonzero(np.diff(rows) != 1) + 1)
                print(f"rep {rep} handle {i} {d}: rel {diff.max() / scale:.2e}; {rows.size} bad rows in {len(runs)} runs:",
                      [(int(r[0]), int(r.size)) for r in runs[:12]], flush=True)
                r0 = rows[0]
                print("    row", r0, "got/ref:", np.round(out[i][d][r0, :4] / ref[i][d][r0, :4], 4), "diff", diff[r0, :4])
print("bad results:", nbad, "of", reps * nthreads * len(dirs))
