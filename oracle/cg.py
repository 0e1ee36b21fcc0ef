"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's conjugate-gradient script
examples/iterative_solver/grm_solve_cg.jl:109-133 (GRM_vec :75-85), the product taken from the
oracle's dgemm_compressed (oracle.c) so that nothing of the product under test is involved.
`lam` adds lambda I to the operator (the script solves the plain G x = b, lam = 0)."""
from __future__ import annotations

import numpy as np

from . import cport


def grm_vec(plink, snps, indiv, f, v, centered=True):
    """GRM_vec: Zv = dgemm_compressed('T', v); Gv = dgemm_compressed('N', Zv)  (:81-82)."""
    zv = cport.dgemm_compressed("T", plink, snps, indiv, f, v.reshape(-1, 1), centered=centered)
    return cport.dgemm_compressed("N", plink, snps, indiv, f, zv, centered=centered)[:, 0]


def cg_solve(plink, snps, indiv, f, b, x0, lam=0.0, tol=1e-2, max_iter=1000, centered=True):
    """Returns (x, iterations, ||r||); statement order as the reference loop."""
    x = np.array(x0, dtype=np.float64, copy=True)
    gx = grm_vec(plink, snps, indiv, f, x, centered) + lam * x       # :113
    r = b - gx                                                        # :114
    p = r.copy()                                                      # :115
    it = 0
    for it in range(max_iter + 1):
        norm_old = np.linalg.norm(r)                                  # :118
        if it == max_iter or not norm_old >= tol:                     # :119-121
            break
        gp = grm_vec(plink, snps, indiv, f, p, centered) + lam * p    # :122
        alpha = norm_old ** 2 / float(p @ gp)                         # :123
        x += alpha * p                                                # :126
        r -= alpha * gp                                               # :127
        beta = np.linalg.norm(r) ** 2 / norm_old ** 2                 # :128
        p = r + beta * p                                              # :129
    return x, it, float(np.linalg.norm(r))
