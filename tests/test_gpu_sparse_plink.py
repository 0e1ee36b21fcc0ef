"""sparse_times_plink / dgemm_plink on the GPU against the UNMODIFIED
reference CPU library (oracle/_ref, sparse_times_plink of src/miraculix/5codesAPI.c:131-157) and a
dense numpy product; fp64, tolerance 1e-10 (sums of at most a few hundred terms: 1e-13 in practice).
dgemm_plink is a BUG() stub in the reference, so it is checked against the oracle's product."""
import ctypes as C

import numpy as np
import pytest

from oracle import cport, ref, simulate

pytestmark = pytest.mark.gpu


def csr(dense):
    rowptr, cols, vals = [0], [], []
    for row in dense:
        nz = np.nonzero(row)[0]
        cols += list(nz)
        vals += list(row[nz])
        rowptr.append(len(cols))
    return (np.array(rowptr, dtype=np.int32), np.array(cols, dtype=np.int32),
            np.array(vals, dtype=np.float64))


def ref_sparse(tc, plink, plink_t, snps, indiv, rp, ci, va, ldc, width):
    L = ref.lib()
    ref.set_options(cores=2, centered=False)
    L.sparse_times_plink.restype = None
    L.sparse_times_plink.argtypes = [C.c_char_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                     C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    out = np.full((ldc, width), 7.0, order="F")
    L.sparse_times_plink(b"N", tc, plink.ctypes.data, plink_t.ctypes.data, snps, indiv, len(rp) - 1,
                         rp.ctypes.data, ci.ctypes.data, va.ctypes.data, out.ctypes.data, ldc)
    return out


@pytest.mark.parametrize("snps,indiv,n_idx,density", [(77, 53, 9, 0.2), (1003, 2051, 40, 0.05),
                                                      (5000, 640, 17, 0.01), (64, 4096, 3, 0.5)])
def test_sparse_times_plink_matches_reference(gpu_lib, snps, indiv, n_idx, density):
    from miraculix_b200 import api
    plink, Z, f = simulate.simulate(snps, indiv, seed=snps + indiv, missing=0.03)
    plink_t = cport.transpose_2bit(plink, snps, indiv)
    rng = np.random.default_rng(n_idx)
    for tc, ncols, width, Zm in (("N", snps, indiv, Z.T.astype(float)), ("T", indiv, snps, Z.astype(float))):
        dense = np.where(rng.random((n_idx, ncols)) < density, rng.standard_normal((n_idx, ncols)), 0.0)
        dense[min(3, n_idx - 1)] = 0.0                      # an empty sparse row
        rp, ci, va = csr(dense)
        ldc = n_idx + 2
        got = api.sparse_times_plink(tc, plink, plink_t, snps, indiv, rp, ci, va, ldc=ldc)
        want = dense @ Zm
        scale = max(np.abs(want).max(), 1e-300)
        assert np.abs(got[:n_idx] - want).max() <= 1e-10 * scale
        assert np.all(got[n_idx:] == 0.0)                   # the reference zeroes the whole Ldc block
        if ref.available():
            r = ref_sparse(tc.encode(), plink, plink_t, snps, indiv, rp, ci, va, ldc, width)
            assert np.abs(got - r).max() <= 1e-10 * scale


def test_dgemm_plink_one_shot(gpu_lib):
    from miraculix_b200 import api
    snps, indiv, n = 1500, 320, 4
    plink, Z, f = simulate.simulate(snps, indiv, seed=8, missing=0.02)
    rng = np.random.default_rng(9)
    for tr, k in (("N", snps), ("T", indiv)):
        B = np.asfortranarray(rng.standard_normal((k, n)))
        for freq, centered in ((None, False), (f, True)):
            got = api.dgemm_plink(tr, plink, snps, indiv, B, freq)
            want = cport.dgemm_compressed(tr, plink, snps, indiv, f, B, centered)
            assert np.abs(got - want).max() <= 1e-10 * np.abs(want).max(), (tr, centered)
    # the process-global centring option is left as it was
    h = api.init_compressed(plink, snps, indiv, f, 1)
    try:
        B = np.asfortranarray(rng.standard_normal((snps, 1)))
        want = cport.dgemm_compressed("N", plink, snps, indiv, f, B, True)
        assert np.abs(api.dgemm_compressed("N", h, B, snps, indiv) - want).max() <= 1e-10 * np.abs(want).max()
    finally:
        api.free_compressed(h)
