"""One rank's panel of the 2-D partitioned GRM on ONE GPU: one CTA per tile vs CTA pairs, timing
and bit-identity.  usage: part_check.py snps indiv nparts [part]"""
import ctypes as C
import sys

import torch

from miraculix_b200 import api

snps, indiv, nparts = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
part = int(sys.argv[4]) if len(sys.argv) > 4 else 0
lib = api.load_library()
dev = torch.device("cuda:0")
bs = (indiv + 3) // 4
bed = torch.empty(snps * bs, dtype=torch.uint8, device=dev)
api._check(lib.mb200_simulate_plink_device(C.c_void_p(bed.data_ptr()), snps, indiv, 4242), "simulate")
h = C.c_void_p(None)
api._check(lib.mb200_plink2compressed_device(C.c_void_p(bed.data_ptr()), snps, indiv, None, 1, C.byref(h)), "p2c")
del bed
torch.cuda.empty_cache()
lay = api.get_layout(h)
kcols = lay.pitch_indiv_major * 4
f = torch.zeros(kcols, dtype=torch.float64, device=dev)
f[:snps].copy_(torch.frombuffer(bytearray(api.get_compressed_freq(h, snps).tobytes()), dtype=torch.float64))
lrows = lib.mb200_crossprod_partition_rows(indiv, part, nparts)
panels = {}
import os
MODES = tuple(int(m) for m in os.environ.get('PART_MODES', '1,2').split(','))
for mode in MODES:
    lib.mb200_set_crossprod_mode(mode)
    panel = torch.empty(max(1, lrows) * indiv, dtype=torch.float64, device=dev)
    ts = []
    for it in range(4):
        api._check(lib.mb200_crossprod_partition_device(C.c_void_p(lay.d_indiv_major), kcols, indiv, 0, 1,
                                                        C.c_void_p(f.data_ptr()), 1, part, nparts,
                                                        C.c_void_p(panel.data_ptr())), "partition")
        t = api.crossprod_last_timing()
        ts.append(t["gemm_ms"] + t["epilogue_ms"])
    ops = t["executed_ops"]
    best = min(ts[1:])
    print(f"mode {mode}: part {part}/{nparts} lrows {lrows}  best {best:.2f} ms  mean {sum(ts[1:])/3:.2f} ms  "
          f"executed {ops/best/1e9:.0f} TOPS  useful-share-of-full {2.0*indiv*indiv*snps/nparts/best/1e9:.0f} TOPS/GPU",
          flush=True)
    panels[mode] = panel
torch.cuda.synchronize()
if len(panels) == 2:
    print("bit-identical:", bool(torch.equal(panels[1], panels[2])))
lib.mb200_set_crossprod_mode(2)
