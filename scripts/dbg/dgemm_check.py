import numpy as np, sys
sys.path.insert(0, '.')
from miraculix_b200 import api
from oracle import cport, simulate, cg as ocg
for snps, indiv in [(2048, 512), (3000, 400), (2048, 500), (2000, 512)]:
    plink, Z, f = simulate.simulate(snps, indiv, seed=17 + snps)
    rng = np.random.default_rng(1)
    for nc in (False, True):
        api.set_options(use_gpu=True, not_center=nc)
        h = api.init_compressed(plink, snps, indiv, f, 1)
        for tr in "NT":
            B = rng.standard_normal((snps if tr == "N" else indiv, 1))
            got = api.dgemm_compressed(tr, h, B, snps, indiv)
            want = cport.dgemm_compressed(tr, plink, snps, indiv, f, B, centered=not nc)
            print(snps, indiv, "not_center", nc, tr, np.abs(got - want).max() / np.abs(want).max())
        b = rng.standard_normal(indiv); x0 = rng.standard_normal(indiv)
        lam = 0.05 * snps
        for iters in (1, 2, 4, 8, 12):
            x, it, res = api.cg_solve(h, b, x0, lam=lam, tol=0.0, max_iter=iters)
            xo, ito, reso = ocg.cg_solve(plink, snps, indiv, f, b, x0, lam=lam, tol=0.0, max_iter=iters, centered=not nc)
            print("   cg iters", iters, "relerr", np.abs(x - xo).max() / np.abs(xo).max(), res, reso)
        api.free_compressed(h)
