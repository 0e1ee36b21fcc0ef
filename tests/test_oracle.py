"""CPU tests: pin the oracle (oracle/oracle.c) against the reference's golden vectors.

Golden fixtures (tests/golden/, made by make_golden.py from the reference itself):
  * plink2twobit_lut.npy  = d_conversion_table of src/cuda/snp_multiply_cuda.h:202
  * dgemm_*.npz           = outputs of the unmodified reference CPU 5codes library
Procedures re-expressed from the reference's tests: tests/dgemm_compressed/test.jl:97-104
(dense product on the decompressed matrix), tests/crossproduct/test_grm.jl:127-139 (raw counts
exact, centred GRM vs BLAS).
"""
import glob
import os

import numpy as np
import pytest

from oracle import cport, ref, simulate

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-10  # north-star tolerance for fp64 outputs: max|dC| / max|C|


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def test_lut_matches_reference_table():
    gold = np.load(os.path.join(GOLDEN, "plink2twobit_lut.npy"))
    assert np.array_equal(cport.plink2twobit_lut(), gold)
    # spot values spelled out in the reference table
    assert gold[0b00000000] == 0 and gold[0b00000010] == 1 and gold[0b00000011] == 2
    assert gold[0b00000001] == 0xFF and gold[0b11111111] == 0xAA


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "dgemm_*.npz"))))
def test_oracle_dgemm_matches_reference_outputs(path):
    g = np.load(path)
    centered = path.endswith("_c1.npz")
    snps, indiv = int(g["snps"]), int(g["indiv"])
    for tr in "NT":
        C = cport.dgemm_compressed(tr, g["plink"], snps, indiv, g["f"], g[f"B_{tr}"], centered)
        assert relerr(C, g[f"C_{tr}"]) <= TOL, (path, tr)


def test_oracle_dgemm_matches_dense_numpy():
    snps, indiv = 317, 129
    plink, Z, f = simulate.simulate(snps, indiv, seed=7, missing=0.02)
    rng = np.random.default_rng(8)
    Zc = Z.astype(np.float64) - 2.0 * f[None, :]
    B = rng.standard_normal((snps, 3))
    assert relerr(cport.dgemm_compressed("N", plink, snps, indiv, f, B), Zc @ B) < 1e-13
    B = rng.standard_normal((indiv, 4))
    assert relerr(cport.dgemm_compressed("T", plink, snps, indiv, f, B), Zc.T @ B) < 1e-13
    assert relerr(cport.dgemm_compressed("T", plink, snps, indiv, f, B, centered=False),
                  Z.T.astype(np.float64) @ B) < 1e-13


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/ref_miraculix.so not built")
def test_oracle_matches_live_reference_library():
    snps, indiv = 2051, 333
    plink, Z, f = simulate.simulate(snps, indiv, seed=11, missing=0.01)
    ref.set_options(cores=4, centered=True)
    h = ref.RefCompressed(plink, snps, indiv, f, 7)
    rng = np.random.default_rng(12)
    for tr, k in (("N", snps), ("T", indiv)):
        B = np.asfortranarray(rng.standard_normal((k, 7)))
        assert relerr(cport.dgemm_compressed(tr, plink, snps, indiv, f, B), h.dgemm(tr, B)) <= TOL


@pytest.mark.parametrize("snps,indiv", [(1, 1), (5, 3), (64, 64), (1003, 203), (513, 1025)])
def test_repack_restatement(snps, indiv):
    plink, Z, _ = simulate.simulate(snps, indiv, seed=snps + indiv, missing=0.05)
    sm, im = cport.repack(plink, snps, indiv)
    assert sm.shape[1] % 128 == 0 and im.shape[1] % 128 == 0
    # round trip: decode(device layout) == decode(bed), element-wise
    assert np.array_equal(cport.unpack_device(im, snps), Z)
    assert np.array_equal(cport.unpack_device(sm, indiv), Z.T)
    # padding is zero
    assert not cport.unpack_device(im, im.shape[1] * 4)[:, snps:].any()
    assert not cport.unpack_device(sm, sm.shape[1] * 4)[:, indiv:].any()
    # transposing twice gives the original (tests/dgemm_compressed/test.jl:79-83)
    t = cport.transpose_2bit(plink, snps, indiv)
    assert np.array_equal(cport.transpose_2bit(t, indiv, snps), plink)
    assert np.array_equal(cport.allele_counts(plink, snps, indiv), Z.sum(axis=0))


@pytest.mark.parametrize("snps,indiv", [(953, 752), (200, 37)])
def test_crossprod_exact(snps, indiv):
    """test_grm.jl:127-131: raw counts must be exact."""
    _, Z, _ = simulate.simulate(snps, indiv, seed=3)
    tb = simulate.pack_2bit(Z)
    M = cport.crossprod(tb, snps, indiv)
    want = Z.astype(np.int64) @ Z.T.astype(np.int64)
    assert np.array_equal(M, want.astype(np.float64))
    assert np.array_equal(M, M.T)


def test_crossprod_plink_format_quirk():
    """With is_plink_format a byte holding a missing code becomes 0xff = four 3s
    (src/cuda/snp_multiply_cuda.h:202), exactly as the reference computes it."""
    snps, indiv = 64, 9
    plink, Z, _ = simulate.simulate(snps, indiv, seed=5, missing=0.1)
    codes = simulate.unpack_2bit(plink, indiv).T  # (indiv, snps) PLINK codes
    x = simulate.pack_2bit(codes)                  # individual-major PLINK-coded
    lut = np.load(os.path.join(GOLDEN, "plink2twobit_lut.npy"))
    vals = simulate.unpack_2bit(lut[x], x.shape[1] * 4).astype(np.int64)
    assert np.array_equal(cport.crossprod(x, snps, indiv, True), (vals @ vals.T).astype(np.float64))


def test_grm_epilogue_matches_formula():
    """docs/grm.md:3-10: G = P Z Z^T P / (2 sum f(1-f)), P = I - 11^T/n."""
    snps, indiv = 400, 57
    _, Z, f = simulate.simulate(snps, indiv, seed=9)
    M = (Z.astype(np.int64) @ Z.T.astype(np.int64)).astype(np.float64)
    G = cport.grm_epilogue(M, f, True)
    P = np.eye(indiv) - np.ones((indiv, indiv)) / indiv
    want = P @ M @ P / (2.0 * np.sum(f * (1.0 - f)))
    assert relerr(G, want) < 1e-12
    # centred via allele freq of the data: equals (Z-2f)(Z-2f)^T scaled (test_grm.jl:133-139)
    Zc = Z - 2.0 * f[None, :]
    assert relerr(G, Zc @ Zc.T / (2.0 * np.sum(f * (1.0 - f)))) < 1e-10


def test_ld_epilogue_is_correlation():
    snps, indiv = 40, 500
    _, Z, f = simulate.simulate(snps, indiv, seed=10)
    M = (Z.T.astype(np.int64) @ Z.astype(np.int64)).astype(np.float64)
    R = cport.ld_epilogue(M, indiv, f)
    want = np.corrcoef(Z.T.astype(np.float64))
    assert relerr(R, want) < 1e-10


def test_oracle_cg_restatement_solves_the_dense_system():
    """oracle/cg.py (the reference script's loop, grm_solve_cg.jl:109-133) against a dense numpy
    solve of (Zc Zc^T + lam I) x = b, and against the live reference library as the product."""
    from oracle import cg as ocg
    snps, indiv = 700, 90
    plink, Z, f = simulate.simulate(snps, indiv, seed=9, missing=0.01)
    rng = np.random.default_rng(10)
    b, x0 = rng.standard_normal(indiv), rng.standard_normal(indiv)
    lam = 0.05 * snps
    Zc = Z.astype(np.float64) - 2.0 * f[None, :]            # indiv x snps
    G = Zc @ Zc.T + lam * np.eye(indiv)
    x, it, res = ocg.cg_solve(plink, snps, indiv, f, b, x0, lam=lam, tol=1e-9, max_iter=400)
    assert it < 400 and res < 1e-9
    assert np.abs(x - np.linalg.solve(G, b)).max() <= 1e-9 * np.abs(x).max()
    v = rng.standard_normal(indiv)
    assert np.abs(ocg.grm_vec(plink, snps, indiv, f, v) - Zc @ (Zc.T @ v)).max() <= 1e-9 * snps


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref/ref_miraculix.so not built")
@pytest.mark.parametrize("seed", range(8))
def test_oracle_matches_live_reference_on_random_shapes(seed):
    """Seeded random ragged shapes, missing codes, centred or not, an EXTERNAL f that is not the
    data's allele frequency (the GPU mode of the reference requires caller-supplied f,
    5codesChar.cc:192-193), 1..17 columns: the restatement against the reference CPU library."""
    rng = np.random.default_rng(1000 + seed)
    snps, indiv = int(rng.integers(1, 3000)), int(rng.integers(1, 1200))
    ncol = int(rng.integers(1, 18))
    centered = bool(seed % 2 == 0)
    plink, Z, f = simulate.simulate(snps, indiv, seed=seed, missing=float(rng.choice([0.0, 0.01, 0.2])))
    if seed % 3 == 0:
        f = rng.uniform(0.0, 1.0, snps)
    ref.set_options(cores=2, centered=centered)
    h = ref.RefCompressed(plink, snps, indiv, f, ncol)
    try:
        for tr, k in (("N", snps), ("T", indiv)):
            B = np.asfortranarray(rng.standard_normal((k, ncol)) * 10.0 ** float(rng.integers(-3, 4)))
            got = cport.dgemm_compressed(tr, plink, snps, indiv, f, B, centered)
            assert relerr(got, h.dgemm(tr, B)) <= TOL, (snps, indiv, ncol, tr, centered)
    finally:
        h.free()
        ref.set_options(cores=4, centered=True)
