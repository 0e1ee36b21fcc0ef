#!/usr/bin/env python
"""PRINT_LEVEL=1 phase timing of the host-pointer GRM (configs[1] size by default)."""
import ctypes as C, os, sys, time
os.environ["PRINT_LEVEL"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from miraculix_b200 import api, build
snps, indiv = (int(x) for x in (sys.argv[1:3] or (50_000, 10_000)))
pinned = (sys.argv[3:4] or ["1"])[0] == "1"
build.build(); lib = api.load_library()
h = api.simulate_compressed(snps, indiv, 4242, 1)
lay = api.get_layout(h)
kcols = lay.pitch_indiv_major * 4
class _H: pass
nbytes = indiv * lay.pitch_indiv_major
hold = type("H", (), {"__cuda_array_interface__": {"shape": (nbytes,), "typestr": "|u1", "data": (int(lay.d_indiv_major), False), "version": 3, "strides": None}})()
hx = torch.as_tensor(hold, device="cuda").cpu().numpy().reshape(indiv, -1)
f = np.zeros(kcols); f[:snps] = api.get_compressed_freq(h, snps)
api.free_compressed(h)
if pinned:
    t = torch.empty(indiv * indiv, dtype=torch.float64).pin_memory(); ans = t.numpy().reshape(indiv, indiv)
    tx = torch.from_numpy(hx.copy()).pin_memory(); hx = tx.numpy()
else:
    ans = np.zeros((indiv, indiv))
for it in range(3):
    t0 = time.perf_counter()
    api._check(lib.mb200_grm(C.c_void_p(hx.ctypes.data), kcols, indiv, C.c_void_p(ans.ctypes.data), False, C.c_void_p(f.ctypes.data), True), "grm")
    print("call %d: %.2f ms" % (it, (time.perf_counter() - t0) * 1e3), flush=True)
