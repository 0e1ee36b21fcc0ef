"""CTA-pair GRM kernel (crossprod mode 2) against mode 1: bit-exact counts and GRM on ragged shapes,
K slabs and the K-split tail; then timing at configs[1]."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from miraculix_b200 import api
from oracle import simulate
lib = api.load_library()
ok = True
for snps, indiv, slab in [(953, 752, None), (3001, 1025, None), (5000, 600, "0.0005"), (9300, 3000, None), (2000, 130, None), (700, 4000, None)]:
    _, Z, f = simulate.simulate(snps, indiv, seed=snps)
    x = simulate.pack_2bit(Z)
    if slab: os.environ["MB200_SLAB_GB"] = slab
    res = {}
    for mode in (1, 2):
        lib.mb200_set_crossprod_mode(mode)
        M = api.snp_crossprod(x, snps, indiv, is_snpmajor=False)
        G = api.grm(x, snps, indiv, do_scale=True, allele_freq=f)
        res[mode] = (M, G)
    os.environ.pop("MB200_SLAB_GB", None)
    Zi = Z.astype(np.int64)
    exact = Zi @ Zi.T
    e1 = np.array_equal(res[1][0], exact); e2 = np.array_equal(res[2][0], exact)
    g = np.array_equal(res[1][1], res[2][1])
    print(f"{snps}x{indiv} slab={slab}: counts mode1 exact {e1}, mode2 exact {e2}, GRM bit-identical {g}", flush=True)
    ok &= e1 and e2 and g
lib.mb200_set_crossprod_mode(1)
print("PAIR PARITY", "OK" if ok else "FAILED")
