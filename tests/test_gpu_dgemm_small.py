"""GPU tests of the latency path of dgemm_compressed (csrc/dgemm_small.cu): small problems through
the host-pointer C-ABI call run as one fused fp64 kernel (centring + product).  Bar: max|dC|/max|C|
<= 1e-10 against the oracle and the reference's golden outputs (north_star); in fact fp64 rounding,
and bit-for-bit reproducible.  The rest of the GPU suite runs with the path switched off
(tests/conftest.py), so everything here switches it on explicitly.
"""
import ctypes as C
import glob
import os

import numpy as np
import pytest

from oracle import cport, simulate

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-10
DEFAULT_LIMIT = 4.0e7


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.fixture(autouse=True)
def _small_path_on(gpu_lib):
    from miraculix_b200 import api
    api.set_options(use_gpu=True, not_center=False)
    api.set_devices(None)
    gpu_lib.mb200_set_dgemm_mode(1, 0)
    gpu_lib.mb200_set_dgemm_small_limit(DEFAULT_LIMIT)
    yield
    gpu_lib.mb200_set_dgemm_small_limit(0.0)
    gpu_lib.mb200_set_dgemm_mode(1, 0)
    api.set_options(use_gpu=True, not_center=False)


def used_small_path(gpu_lib, h):
    sl = C.c_int(-1)
    return gpu_lib.mb200_dgemm_last_precision(h, C.byref(sl)) == 0 and sl.value == 0


@pytest.mark.parametrize("not_center", [False, True])
@pytest.mark.parametrize("ncol", [1, 2, 3, 4, 5, 8, 10, 13])
@pytest.mark.parametrize("snps,indiv,missing", [(1, 1, 0.0), (5, 3, 0.0), (1003, 203, 0.02), (513, 1025, 0.05),
                                                (10000, 1000, 0.0), (300, 4099, 0.01), (20000, 130, 0.0)])
def test_small_path_matches_oracle(gpu_lib, snps, indiv, missing, ncol, not_center):
    from miraculix_b200 import api
    if snps * indiv * ncol > DEFAULT_LIMIT:
        pytest.skip("above the work limit of the small path")
    plink, Z, f = simulate.simulate(snps, indiv, seed=snps + indiv + ncol, missing=missing)
    rng = np.random.default_rng(ncol)
    api.set_options(use_gpu=True, not_center=not_center)
    h = api.init_compressed(plink, snps, indiv, f, ncol)
    try:
        for tr, k in (("N", snps), ("T", indiv)):
            B = np.asfortranarray(rng.standard_normal((k, ncol)))
            got = api.dgemm_compressed(tr, h, B, snps, indiv)
            assert used_small_path(gpu_lib, h)
            want = cport.dgemm_compressed(tr, plink, snps, indiv, f, B, not not_center)
            assert relerr(got, want) <= 1e-13, (tr, relerr(got, want))
            again = api.dgemm_compressed(tr, h, B, snps, indiv)
            assert np.array_equal(got, again), "fixed summation order: results must repeat bit for bit"
    finally:
        api.free_compressed(h)


def test_small_path_reference_goldens(gpu_lib):
    """The reference CPU library's own outputs (tests/golden/dgemm_*.npz, made by make_golden.py)."""
    from miraculix_b200 import api
    files = sorted(glob.glob(os.path.join(GOLDEN, "dgemm_*.npz")))
    assert files
    for path in files:
        g = np.load(path)
        centered = path.endswith("_c1.npz")
        snps, indiv = int(g["snps"]), int(g["indiv"])
        assert snps * indiv * g["B_N"].shape[1] <= DEFAULT_LIMIT
        api.set_options(use_gpu=True, not_center=not centered)
        h = api.init_compressed(g["plink"], snps, indiv, g["f"], 16)
        try:
            for tr in "NT":
                Cg = api.dgemm_compressed(tr, h, g[f"B_{tr}"], snps, indiv)
                assert used_small_path(gpu_lib, h)
                assert relerr(Cg, g[f"C_{tr}"]) <= TOL, (os.path.basename(path), tr)
        finally:
            api.free_compressed(h)


def test_small_path_leading_dimensions_limits_and_mixing(gpu_lib):
    """ldb / ldc larger than the panels (padding rows untouched); the work limit, fixed digits and
    mode 0 keep the other kernels; small and tensor-core calls alternate on one handle."""
    from miraculix_b200 import api
    snps, indiv, n = 3000, 500, 4
    plink, Z, f = simulate.simulate(snps, indiv, seed=3, missing=0.01)
    rng = np.random.default_rng(4)
    h = api.init_compressed(plink, snps, indiv, f, n)
    try:
        ldb, ldc = snps + 13, indiv + 5
        Bbig = np.asfortranarray(rng.standard_normal((ldb, n)))
        Cbig = np.full((ldc, n), 777.0, order="F")
        gpu_lib.dgemm_compressed(b"N", h, n, Bbig.ctypes.data, ldb, Cbig.ctypes.data, ldc)
        assert gpu_lib.mb200_last_status() == 0 and used_small_path(gpu_lib, h)
        want_n = cport.dgemm_compressed("N", plink, snps, indiv, f, Bbig[:snps], True)
        assert relerr(Cbig[:indiv], want_n) <= 1e-13 and np.all(Cbig[indiv:] == 777.0)
        ldb, ldc = indiv + 3, snps + 7
        Bt = np.asfortranarray(rng.standard_normal((ldb, n)))
        Ct = np.full((ldc, n), -5.0, order="F")
        gpu_lib.dgemm_compressed(b"T", h, n, Bt.ctypes.data, ldb, Ct.ctypes.data, ldc)
        assert gpu_lib.mb200_last_status() == 0 and used_small_path(gpu_lib, h)
        want_t = cport.dgemm_compressed("T", plink, snps, indiv, f, Bt[:indiv], True)
        assert relerr(Ct[:snps], want_t) <= 1e-13 and np.all(Ct[snps:] == -5.0)
        # n > max_n is still refused, C untouched
        B5 = np.asfortranarray(rng.standard_normal((snps, n + 1)))
        C5 = np.zeros((indiv, n + 1), order="F")
        gpu_lib.dgemm_compressed(b"N", h, n + 1, B5.ctypes.data, snps, C5.ctypes.data, indiv)
        assert gpu_lib.mb200_last_status() != 0 and not C5.any()
        # above the limit -> tensor-core path (digit slices); same answer to 1e-10
        B = np.asfortranarray(Bbig[:snps])
        gpu_lib.mb200_set_dgemm_small_limit(float(snps * indiv * n - 1))
        got = api.dgemm_compressed("N", h, B, snps, indiv)
        sl = C.c_int(0)
        assert gpu_lib.mb200_dgemm_last_precision(h, C.byref(sl)) >= 0 and sl.value in (6, 7, 8)
        assert relerr(got, want_n) <= TOL
        gpu_lib.mb200_set_dgemm_small_limit(DEFAULT_LIMIT)
        got2 = api.dgemm_compressed("N", h, B, snps, indiv)
        assert used_small_path(gpu_lib, h) and relerr(got2, want_n) <= 1e-13
        # fixed digit counts and the fp64 SIMT family are explicit requests for those kernels
        for mode, slices in ((1, 6), (1, 8), (0, 0)):
            gpu_lib.mb200_set_dgemm_mode(mode, slices)
            got = api.dgemm_compressed("N", h, B, snps, indiv)
            assert relerr(got, want_n) <= TOL
            if mode == 1:
                assert gpu_lib.mb200_dgemm_last_precision(h, C.byref(sl)) == 0 and sl.value == slices
        gpu_lib.mb200_set_dgemm_mode(1, 0)
    finally:
        api.free_compressed(h)


def test_small_path_nonfinite_b_propagates(gpu_lib):
    """NaN / Inf in B poison their column as in the reference's fp64 GEMM (0 * NaN = NaN, so every
    row of the column); the other columns are unaffected."""
    from miraculix_b200 import api
    snps, indiv = 700, 64
    plink, Z, f = simulate.simulate(snps, indiv, seed=9)
    api.set_options(use_gpu=True, not_center=True)
    h = api.init_compressed(plink, snps, indiv, f, 3)
    try:
        B = np.asfortranarray(np.random.default_rng(1).standard_normal((snps, 3)))
        B[17, 0] = np.nan
        B[100, 1] = np.inf
        got = api.dgemm_compressed("N", h, B, snps, indiv)
        assert used_small_path(gpu_lib, h)
        assert np.all(np.isnan(got[:, 0])) and not np.isfinite(got[:, 1]).any()
        want = cport.dgemm_compressed("N", plink, snps, indiv, f, B[:, 2:3], False)
        assert relerr(got[:, 2:3], want) <= 1e-13
    finally:
        api.free_compressed(h)


@pytest.mark.parametrize("b_pinned,c_pinned", [(True, True), (True, False), (False, True)])
def test_small_path_page_locked_panels(gpu_lib, b_pinned, c_pinned):
    """Page-locked caller panels are copied directly, pageable ones through the pinned ring, in any
    combination."""
    import torch
    from miraculix_b200 import api
    snps, indiv, n = 4100, 300, 3
    plink, Z, f = simulate.simulate(snps, indiv, seed=21, missing=0.01)
    h = api.init_compressed(plink, snps, indiv, f, n)
    try:
        for tr, k, m in (("N", snps, indiv), ("T", indiv, snps)):
            tb = torch.randn(n, k, dtype=torch.float64)         # row-major (n, k) == column-major k x n
            tc = torch.full((n, m), 9.0, dtype=torch.float64)
            if b_pinned:
                tb = tb.pin_memory()
            if c_pinned:
                tc = tc.pin_memory()
            gpu_lib.dgemm_compressed(tr.encode(), h, n, tb.data_ptr(), k, tc.data_ptr(), m)
            assert gpu_lib.mb200_last_status() == 0 and used_small_path(gpu_lib, h)
            want = cport.dgemm_compressed(tr, plink, snps, indiv, f, tb.numpy().T, True)
            assert relerr(tc.numpy().T, want) <= 1e-13
    finally:
        api.free_compressed(h)
