"""Latency of the host-pointer dgemm_compressed call on small problems: one-kernel path
(dgemm_small.cu) against the tensor-core path, pageable numpy panels (or pinned: argv[1] = pinned),
wall clock per call and the device time between the events around the kernels."""
import ctypes as C, statistics, sys, time
import numpy as np
sys.path.insert(0, ".")
from miraculix_b200 import api
lib = api.load_library()
api.set_options(use_gpu=True, not_center=False)
pinned = len(sys.argv) > 1 and sys.argv[1] == "pinned"
def lat(snps, indiv, n, tr, limit):
    lib.mb200_set_dgemm_small_limit(float(limit))
    h = api.simulate_compressed(snps, indiv, 5, n)
    k, m = (indiv, snps) if tr == "T" else (snps, indiv)
    B = np.asfortranarray(np.random.default_rng(1).standard_normal((k, n)))
    Cout = np.empty((m, n), order="F")
    if pinned:
        import torch
        tb, tc = torch.from_numpy(B.T.copy()).pin_memory(), torch.empty(n, m, dtype=torch.float64).pin_memory()
        pb, pc = tb.data_ptr(), tc.data_ptr()
    else:
        pb, pc = B.ctypes.data, Cout.ctypes.data
    ts, ks = [], []
    global first_call
    ms = C.c_float(0)
    for it in range(150):
        t0 = time.perf_counter()
        lib.dgemm_compressed(tr.encode(), h, n, pb, k, pc, m)
        t1 = time.perf_counter()
        if it == 0:
            first_call = (t1 - t0) * 1e6
        if it >= 30:
            ts.append(t1 - t0)
            lib.mb200_dgemm_last_kernel_ms(h, C.byref(ms))
            ks.append(ms.value)
    assert lib.mb200_last_status() == 0
    lib.free_compressed(C.byref(h))
    return statistics.median(ts) * 1e6, statistics.median(ks) * 1e3
first_call = 0.0
shapes = [(512, 8, 1), (10000, 1000, 1), (10000, 1000, 4), (10000, 1000, 10), (30000, 1000, 1), (30000, 3000, 1),
          (100000, 1000, 1)]
for snps, indiv, n in shapes:
    for tr in "NT":
        (a, ka), (b, kb) = lat(snps, indiv, n, tr, 1e18), lat(snps, indiv, n, tr, 0)
        print(f"first call of the handle (tensor-core path, pageable): {first_call:8.1f} us", end="   ") if not pinned else None
        print(f"{snps:7d} x {indiv:5d} n={n:2d} '{tr}'  work {snps*indiv*n:.1e}  small {a:7.1f} us (kernel {ka:6.1f})"
              f"   tensor-core {b:7.1f} us (main kernel {kb:6.1f})", flush=True)
