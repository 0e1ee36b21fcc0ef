set -x; mkdir -p gpurun_out
timeout 300 python scripts/dbg/chunk_sweep.py 125000 100000 10 2>&1 | tail -8
timeout 300 python scripts/dbg/chunk_sweep.py 250000 100000 10 2>&1 | tail -8
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q --timeout 300 -k "tiled or grm or pageable" 2>&1 | tail -3
PRINT_LEVEL=1 MB200_E2E=1 timeout 300 python - <<'PY' 2>&1 | grep -a "crossproduct\|setup\|enqueued\|drained\|released\|call" | tail -16
import os, sys, time, ctypes as C
sys.path.insert(0, os.getcwd())
import numpy as np, torch
from miraculix_b200 import api, build
lib = api.load_library()
snps, indiv = 50000, 10000
h = api.simulate_compressed(snps, indiv, 4242, 1); lay = api.get_layout(h); kcols = lay.pitch_indiv_major * 4
nbytes = indiv * lay.pitch_indiv_major
hold = type("H", (), {"__cuda_array_interface__": {"shape": (nbytes,), "typestr": "|u1", "data": (int(lay.d_indiv_major), False), "version": 3, "strides": None}})()
hx = torch.as_tensor(hold, device="cuda").cpu().pin_memory().numpy().reshape(indiv, -1)
f = np.zeros(kcols); f[:snps] = api.get_compressed_freq(h, snps); api.free_compressed(h)
t = torch.empty(indiv * indiv, dtype=torch.float64).pin_memory(); ans = t.numpy().reshape(indiv, indiv)
for devs in ([0, 0], [0, 0, 0, 0]):
    api.set_devices(devs)
    for it in range(3):
        t0 = time.perf_counter()
        api._check(lib.mb200_grm(C.c_void_p(hx.ctypes.data), kcols, indiv, C.c_void_p(ans.ctypes.data), False, C.c_void_p(f.ctypes.data), True), "grm")
        print("call", devs, it, "%.2f ms" % ((time.perf_counter() - t0) * 1e3), flush=True)
PY
