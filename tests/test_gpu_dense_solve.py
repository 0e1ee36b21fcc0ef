"""potrs_solve_gpu (reference src/cuda/solve_cuda.cu:69-278, Julia binding
solve.jl:139-180) against numpy: fp64, tolerance 1e-10 relative on well-conditioned SPD systems,
log-determinant to 1e-12 relative; a non-SPD matrix must report status 1, not garbage."""
import numpy as np
import pytest

from oracle import simulate

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,nrhs", [(1, 1), (7, 2), (500, 3), (2048, 10)])
def test_dense_solve_matches_numpy(gpu_lib, n, nrhs):
    from miraculix_b200 import api
    rng = np.random.default_rng(n)
    A = rng.standard_normal((n, n + 5))
    M = A @ A.T + n * np.eye(n)
    B = rng.standard_normal((n, nrhs))
    X, logdet = api.dense_solve(M, B, calc_logdet=True)
    want = np.linalg.solve(M, B)
    assert np.abs(X - want).max() <= 1e-10 * np.abs(want).max()
    assert abs(logdet - np.linalg.slogdet(M)[1]) <= 1e-12 * abs(np.linalg.slogdet(M)[1]) + 1e-12
    X2 = api.dense_solve(M, B[:, 0])           # vector right-hand side, no logdet, managed memory
    assert X2.shape == (n,) and np.abs(X2 - want[:, 0]).max() <= 1e-10 * np.abs(want).max()
    X3 = api.dense_solve(M, B, oversubscribe=True)
    assert np.array_equal(X3, X)


def test_dense_solve_rejects_indefinite(gpu_lib):
    from miraculix_b200 import api
    M = np.eye(6)
    M[3, 3] = -1.0
    with pytest.raises(api.MiraculixError, match="minor 4"):
        api.dense_solve(M, np.ones(6))


def test_gblup_chain_grm_then_solve(gpu_lib):
    """examples/gblup/calculate_gblup.jl:138-181 in miniature: G = grm(Z), then (G + I) x = y."""
    from miraculix_b200 import api
    snps, indiv = 3000, 257
    plink, Z, f = simulate.simulate(snps, indiv, seed=4)
    x2 = simulate.pack_2bit(Z)
    G = api.grm(x2, snps, indiv, do_scale=True, allele_freq=f)
    y = np.random.default_rng(5).standard_normal(indiv)
    got = api.dense_solve(G + np.eye(indiv), y)
    Zc = Z.astype(np.float64) - Z.mean(axis=0, keepdims=True)
    Gn = Zc @ Zc.T / (2.0 * np.sum(f * (1.0 - f)))
    want = np.linalg.solve(Gn + np.eye(indiv), y)
    assert np.abs(got - want).max() <= 1e-9 * np.abs(want).max()
