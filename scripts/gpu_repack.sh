set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dgemm.py tests/test_bedfile.py -m gpu -x -q -k "repack or bed" 2>&1 | tail -4
python - <<'PY'
import ctypes as C, time, torch
from miraculix_b200 import api
lib = api.load_library()
snps, indiv = 1_000_000, 100_000
bs = (indiv + 3)//4
bed = torch.empty(snps*bs, dtype=torch.uint8, device="cuda")
api._check(lib.mb200_simulate_plink_device(C.c_void_p(bed.data_ptr()), snps, indiv, 1), "sim")
for rep in range(3):
    h = C.c_void_p(None)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    api._check(lib.mb200_plink2compressed_device(C.c_void_p(bed.data_ptr()), snps, indiv, None, 1, C.byref(h)), "p2c")
    torch.cuda.synchronize(); t1 = time.perf_counter()
    lib.free_compressed(C.byref(h))
    print("plink2compressed_device total %.1f ms (alloc+memset+repack+rowsum maxima)" % ((t1-t0)*1e3))
PY
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:repack -c 4 python bench.py --steps 1 --warmup 1 --no-cpu --no-grm --no-pcg 2>&1 | grep -E "repack|duration|dram__" | head -20
