"""Handle creation at configs[2] on the device: total time of mb200_plink2compressed_device (3 repeats)."""
import ctypes as C, time, sys, torch
from miraculix_b200 import api
lib = api.load_library()
snps, indiv = 1_000_000, 100_000
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
bs = (indiv + 3)//4
bed = torch.empty(snps*bs, dtype=torch.uint8, device="cuda")
api._check(lib.mb200_simulate_plink_device(C.c_void_p(bed.data_ptr()), snps, indiv, 1), "sim")
for rep in range(reps):
    h = C.c_void_p(None)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    api._check(lib.mb200_plink2compressed_device(C.c_void_p(bed.data_ptr()), snps, indiv, None, 1, C.byref(h)), "p2c")
    torch.cuda.synchronize(); t1 = time.perf_counter()
    lib.free_compressed(C.byref(h))
    print("plink2compressed_device total %.1f ms (alloc+repack+rowsum maxima)" % ((t1-t0)*1e3))
