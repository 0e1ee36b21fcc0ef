set -x; mkdir -p gpurun_out
timeout 1800 python -m pytest tests/test_gpu_multi.py tests/test_gpu_sharded_nccl.py -q --timeout 300 2>&1 | tail -6 > gpurun_out/r2r_multi_2gpu.txt; cat gpurun_out/r2r_multi_2gpu.txt
PRINT_LEVEL=1 timeout 300 python - <<'PY' 2>&1 | grep -a "crossproduct\|setup\|enqueued\|drained\|released\|call" | sed 's/\x1b\[[0-9;]*m//g' | tail -34
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
for devs in ([0], [0, 1]):
    api.set_devices(devs if len(devs) > 1 else None)
    for it in range(3):
        t0 = time.perf_counter()
        api._check(lib.mb200_grm(C.c_void_p(hx.ctypes.data), kcols, indiv, C.c_void_p(ans.ctypes.data), False, C.c_void_p(f.ctypes.data), True), "grm")
        print("call", devs, it, "%.2f ms" % ((time.perf_counter() - t0) * 1e3), flush=True)
PY
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2r_bench_n2.json 2> gpurun_out/r2r_bench_n2.err; echo rc=$?; tail -3 gpurun_out/r2r_bench_n2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2r_bench_n2.json'))
e=d['e2e']
print('value', d['value'], 'ms/step', d['ms_per_step'], 'frac', d['roofline']['frac'])
print('e2e', e['ms_per_step'], 'n', e['n_ms'], 't', e['t_ms'], 'pageable', e['pageable']['ms_per_step'], e['parity']['ok'], e['host_link'])
print('pcg', d['pcg']['ms_per_iteration'], d['pcg']['e2e']['ms_per_iteration'])
print('grm', d['grm']['ms'], json.dumps(d['grm']['e2e'])[:500])
PY
