"""GPU parity tests for plink2compressed (repack) and dgemm_compressed, through the C-ABI.

Bars (BASELINE.json north_star): repack bit-exact; fp64 dgemm_compressed max|dC|/max|C| <= 1e-10
against the reference's own CPU implementation (golden outputs in tests/golden/) and against the
oracle on seeded inputs.  Shapes follow the reference's tests: tests/dgemm_compressed/test.jl
(ncol=10, randn B), test_5codesapi.f90 / test_5codesapi_t.f90 (B = -(10 i + j), -(1000 i + j)).
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


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.fixture(autouse=True)
def _centered(gpu_lib):
    from miraculix_b200 import api
    api.set_options(use_gpu=True, not_center=False)
    yield


@pytest.mark.parametrize("snps,indiv,missing", [(1, 1, 0.0), (5, 3, 0.0), (256, 256, 0.0),
                                                (1003, 203, 0.02), (513, 1025, 0.05),
                                                (10000, 1000, 0.0), (300, 4099, 0.01),
                                                # whole-word bed rows: the 512 x 512 in-place
                                                # transpose kernel, ragged in both directions
                                                (1003, 208, 0.02), (5000, 1037, 0.01),
                                                (700, 8208, 0.0), (1537, 16, 0.05), (512, 512, 0.0)])
def test_repack_bit_exact(gpu_lib, snps, indiv, missing):
    from miraculix_b200 import api
    plink, Z, f = simulate.simulate(snps, indiv, seed=snps * 7 + indiv, missing=missing)
    h = api.init_compressed(plink, snps, indiv, f, 1)
    try:
        sm, im, cnt = api.download_layout(h)
        sm_o, im_o = cport.repack(plink, snps, indiv)
        assert sm.shape == sm_o.shape and im.shape == im_o.shape
        assert np.array_equal(sm, sm_o)
        assert np.array_equal(im, im_o)
        assert np.array_equal(cnt.astype(np.int64), cport.allele_counts(plink, snps, indiv))
        # round trip against the bed itself
        assert np.array_equal(cport.unpack_device(im, snps), Z)
        assert np.array_equal(api.get_compressed_freq(h, snps), f)
    finally:
        api.free_compressed(h)
    assert not h.value  # free_compressed nulls the caller's pointer


MODES = [(0, 7), (1, 7), (1, 6), (1, 8), (1, 0)]  # (kernel family, digits; 0 = adaptive)


@pytest.mark.parametrize("mode,slices", MODES)
@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "dgemm_*.npz"))))
def test_dgemm_matches_reference_golden(gpu_lib, path, mode, slices):
    """Outputs of the unmodified reference CPU 5codes library on the same inputs."""
    from miraculix_b200 import api
    g = np.load(path)
    centered = path.endswith("_c1.npz")
    snps, indiv = int(g["snps"]), int(g["indiv"])
    api.set_options(use_gpu=True, not_center=not centered)
    gpu_lib.mb200_set_dgemm_mode(mode, slices)
    h = api.init_compressed(g["plink"], snps, indiv, g["f"], 16)
    try:
        for tr in "NT":
            Cg = api.dgemm_compressed(tr, h, g[f"B_{tr}"], snps, indiv)
            assert relerr(Cg, g[f"C_{tr}"]) <= TOL, (os.path.basename(path), tr, mode)
    finally:
        api.free_compressed(h)
        gpu_lib.mb200_set_dgemm_mode(1, 0)


@pytest.mark.parametrize("mode,slices", MODES)
@pytest.mark.parametrize("snps,indiv,ncol", [(10000, 1000, 1), (1003, 203, 10), (2051, 777, 3),
                                             (512, 128, 16), (4097, 2050, 7)])
def test_dgemm_matches_oracle(gpu_lib, snps, indiv, ncol, mode, slices):
    from miraculix_b200 import api
    plink, Z, f = simulate.simulate(snps, indiv, seed=snps + ncol, missing=0.01)
    rng = np.random.default_rng(ncol)
    gpu_lib.mb200_set_dgemm_mode(mode, slices)
    h = api.init_compressed(plink, snps, indiv, f, ncol)
    try:
        for tr, k in (("N", snps), ("T", indiv)):
            B = np.asfortranarray(rng.standard_normal((k, ncol)))
            want = cport.dgemm_compressed(tr, plink, snps, indiv, f, B, True)
            assert relerr(api.dgemm_compressed(tr, h, B, snps, indiv), want) <= TOL, (tr, mode)
    finally:
        api.free_compressed(h)
        gpu_lib.mb200_set_dgemm_mode(1, 0)


@pytest.mark.parametrize("cluster", [0, 1])
@pytest.mark.parametrize("slices", [0, 6, 8])
@pytest.mark.parametrize("snps,indiv,ncol", [(3000, 700, 50), (1200, 2100, 27), (6000, 300, 14)])
def test_dgemm_wide_b_column_groups(gpu_lib, snps, indiv, ncol, slices, cluster):
    """More columns than one MMA's digit rows hold (13 at 6 digits, 10 at 8): the column groups run
    in ONE launch whose units share the packed tiles (examples/gblup/calculate_gblup.jl:89-93 calls
    with n + p = 50 columns).  Every column against the oracle; cluster = 1: thread-block clusters with
    the packed tiles multicast (2 or 4 groups)."""
    from miraculix_b200 import api
    plink, Z, f = simulate.simulate(snps, indiv, seed=snps + ncol, missing=0.01)
    rng = np.random.default_rng(ncol)
    gpu_lib.mb200_set_dgemm_mode(1, slices)
    gpu_lib.mb200_set_dgemm_wide_cluster(cluster)
    h = api.init_compressed(plink, snps, indiv, f, ncol)
    try:
        for tr, k in (("N", snps), ("T", indiv)):
            B = np.asfortranarray(rng.standard_normal((k, ncol)) * np.logspace(-3, 3, ncol))
            want = cport.dgemm_compressed(tr, plink, snps, indiv, f, B, True)
            got = api.dgemm_compressed(tr, h, B, snps, indiv)
            for c in range(ncol):
                assert relerr(got[:, c], want[:, c]) <= TOL, (tr, c)
    finally:
        api.free_compressed(h)
        gpu_lib.mb200_set_dgemm_mode(1, 0)
        gpu_lib.mb200_set_dgemm_wide_cluster(0)


def test_dgemm_adaptive_precision(gpu_lib):
    """Default digit policy on a problem large enough to take the 6-digit path: the verified
    product meets the 1e-10 bar (precision code 1); a product that cancels to ~0 (centred 'T'
    times the ones vector: Z^T 1 - 2 f indiv = 0) fails the bound and is redone with 8 digits
    (code 2), still matching the oracle to 1e-10 of the uncancelled magnitude."""
    from miraculix_b200 import api
    snps, indiv, ncol = 16384, 20000, 3
    plink, Z, f = simulate.simulate(snps, indiv, seed=99)
    rng = np.random.default_rng(100)
    gpu_lib.mb200_set_dgemm_mode(1, 0)
    h = api.init_compressed(plink, snps, indiv, f, ncol)
    sl = C.c_int(0)
    try:
        for tr, k in (("N", snps), ("T", indiv)):
            B = np.asfortranarray(rng.standard_normal((k, ncol)))
            got = api.dgemm_compressed(tr, h, B, snps, indiv)
            assert gpu_lib.mb200_dgemm_last_precision(h, C.byref(sl)) == 1 and sl.value == 6
            want = cport.dgemm_compressed(tr, plink, snps, indiv, f, B, True)
            assert relerr(got, want) <= TOL, tr
        ones = np.ones((indiv, ncol), order="F")
        got = api.dgemm_compressed("T", h, ones, snps, indiv)
        assert gpu_lib.mb200_dgemm_last_precision(h, C.byref(sl)) == 2 and sl.value == 8
        want = cport.dgemm_compressed("T", plink, snps, indiv, f, ones, True)
        scale = float(Z.sum(axis=0).max())  # magnitude of the two terms that cancel
        assert np.abs(got - want).max() <= TOL * scale
    finally:
        api.free_compressed(h)


def test_dgemm_leading_dimensions_and_max_n(gpu_lib):
    """Ldb/Ldc are honoured (reference GPU path ignores them) and n > max_n is refused."""
    from miraculix_b200 import api
    snps, indiv, n = 700, 300, 4
    plink, Z, f = simulate.simulate(snps, indiv, seed=1)
    rng = np.random.default_rng(2)
    h = api.init_compressed(plink, snps, indiv, f, n)
    try:
        ldb, ldc = snps + 13, indiv + 5
        Bbig = np.asfortranarray(rng.standard_normal((ldb, n)))
        Cbig = np.full((ldc, n), 777.0, order="F")
        gpu_lib.dgemm_compressed(b"n", h, n, Bbig.ctypes.data, ldb, Cbig.ctypes.data, ldc)
        assert gpu_lib.mb200_last_status() == 0
        want = cport.dgemm_compressed("N", plink, snps, indiv, f, Bbig[:snps], True)
        assert relerr(Cbig[:indiv], want) <= TOL
        assert np.all(Cbig[indiv:] == 777.0)  # rows beyond m untouched
        # 'y' is an alias of 'T' (5codesAPI.c:73-77)
        B = np.asfortranarray(rng.standard_normal((indiv, 2)))
        Cy = np.zeros((snps, 2), order="F")
        gpu_lib.dgemm_compressed(b"y", h, 2, B.ctypes.data, indiv, Cy.ctypes.data, snps)
        assert relerr(Cy, cport.dgemm_compressed("T", plink, snps, indiv, f, B, True)) <= TOL
        # n > max_n
        B5 = np.asfortranarray(rng.standard_normal((snps, n + 1)))
        C5 = np.zeros((indiv, n + 1), order="F")
        gpu_lib.dgemm_compressed(b"N", h, n + 1, B5.ctypes.data, snps, C5.ctypes.data, indiv)
        assert gpu_lib.mb200_last_status() != 0 and not C5.any()
    finally:
        api.free_compressed(h)


def test_dgemm_linearity_and_transpose_identity_large(gpu_lib):
    """Size-independent properties at a size the oracle does not finish in seconds:
    linearity in B, and <u, A v> == <A^T u, v> for A = Z - 2 1 f^T."""
    from miraculix_b200 import api
    snps, indiv = 60000, 8000
    plink, _, f = simulate.simulate(snps, indiv, seed=42)
    rng = np.random.default_rng(43)
    h = api.init_compressed(plink, snps, indiv, f, 2)
    try:
        v1, v2 = rng.standard_normal((snps, 1)), rng.standard_normal((snps, 1))
        a1 = api.dgemm_compressed("N", h, v1, snps, indiv)
        a2 = api.dgemm_compressed("N", h, v2, snps, indiv)
        a12 = api.dgemm_compressed("N", h, 2.0 * v1 - 3.0 * v2, snps, indiv)
        assert relerr(a12, 2.0 * a1 - 3.0 * a2) <= 1e-9
        u = rng.standard_normal((indiv, 1))
        atu = api.dgemm_compressed("T", h, u, snps, indiv)
        lhs, rhs = (u.T @ a1).item(), (atu.T @ v1).item()
        assert abs(lhs - rhs) <= 1e-9 * max(abs(lhs), np.linalg.norm(u) * np.linalg.norm(a1))
        # two columns at once == two single calls
        both = api.dgemm_compressed("N", h, np.hstack([v1, v2]), snps, indiv)
        assert relerr(both, np.hstack([a1, a2])) <= 1e-12
    finally:
        api.free_compressed(h)


def test_dgemm_full_size_properties_and_spot_parity(gpu_lib):
    """BASELINE configs[2] at FULL size (1M SNPs x 100k individuals, ncol = 10, device-simulated bed):
    (1) <U, A V> == <A^T U, V> over all ten columns, (2) rows of the 'N' result and of the 'T' result
    for a handful of individuals / SNPs against a plain numpy long-double product on genotypes
    decoded straight from the simulated bed (independent of the repack and of both kernels),
    (3) the stored allele frequencies against the same decode.  Skips on a GPU without ~90 GB free."""
    import torch
    from miraculix_b200 import api
    snps, indiv, n = 1_000_000, 100_000, 10
    bs = (indiv + 3) // 4
    free_b, _ = torch.cuda.mem_get_info()
    if free_b < 90e9:
        pytest.skip("needs ~90 GB of free device memory")
    d_plink = torch.empty(snps * bs, dtype=torch.uint8, device="cuda")
    assert gpu_lib.mb200_simulate_plink_device(d_plink.data_ptr(), snps, indiv, 20261104) == 0
    h = C.c_void_p(None)
    assert gpu_lib.mb200_plink2compressed_device(d_plink.data_ptr(), snps, indiv, None, n,
                                                 C.byref(h)) == 0
    try:
        g = torch.Generator(device="cuda").manual_seed(7)
        V = torch.randn(n, snps, dtype=torch.float64, device="cuda", generator=g)   # col-major snps x n
        U = torch.randn(n, indiv, dtype=torch.float64, device="cuda", generator=g)  # col-major indiv x n
        AV = torch.empty(n, indiv, dtype=torch.float64, device="cuda")
        AtU = torch.empty(n, snps, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        assert gpu_lib.mb200_dgemm_compressed_device(0, h, n, V.data_ptr(), snps, AV.data_ptr(), indiv) == 0
        assert gpu_lib.mb200_dgemm_compressed_device(1, h, n, U.data_ptr(), indiv, AtU.data_ptr(), snps) == 0
        assert gpu_lib.mb200_sync(h) == 0
        lhs = (U * AV).sum(dim=1).cpu().numpy()
        rhs = (AtU * V).sum(dim=1).cpu().numpy()
        scale = (U.norm(dim=1) * AV.norm(dim=1)).cpu().numpy()
        assert (np.abs(lhs - rhs) <= 1e-10 * scale).all()

        f = api.get_compressed_freq(h, snps)
        bed = d_plink.view(snps, bs)
        dec = np.array([0.0, 0.0, 1.0, 2.0])                      # PLINK code -> value (MXinfo.h:143)
        # 'N' rows: individuals i -> z_i[s] for all SNPs, from byte column i // 4 of the bed
        rows_n = [0, 1, 77_777, 99_999]
        Vh = V.cpu().numpy().T.astype(np.longdouble)                # snps x n
        corr = 2.0 * (f.astype(np.longdouble) @ Vh)                 # centring: -2 f^T V
        for i in rows_n:
            col = bed[:, i // 4].cpu().numpy()
            z = dec[(col >> (2 * (i % 4))) & 3].astype(np.longdouble)
            want = (z @ Vh - corr).astype(np.float64)
            got = AV[:, i].cpu().numpy()
            assert np.abs(got - want).max() <= TOL * float(AV.abs().max())
        # 'T' rows: SNPs s -> z[:, s] for all individuals, from bed row s
        Uh = U.cpu().numpy().T.astype(np.longdouble)                # indiv x n
        colsum = Uh.sum(axis=0)
        for sidx in [0, 513, 500_000, 999_999]:
            row = bed[sidx].cpu().numpy()
            z = dec[(row[:, None] >> np.array([0, 2, 4, 6])) & 3].reshape(-1)[:indiv].astype(np.longdouble)
            want = (z @ Uh - 2.0 * np.longdouble(f[sidx]) * colsum).astype(np.float64)
            got = AtU[:, sidx].cpu().numpy()
            assert np.abs(got - want).max() <= TOL * float(AtU.abs().max())
            assert abs(f[sidx] - float(z.sum()) / (2.0 * indiv)) <= 1e-15
    finally:
        api.free_compressed(h)


def test_device_pointer_entry_and_simulator(gpu_lib):
    """mb200_plink2compressed_device + mb200_dgemm_compressed_device on torch device buffers; the
    on-device simulator output is a valid bed (no missing codes, zero pad bits) and is
    deterministic in (seed)."""
    import torch
    from miraculix_b200 import api
    snps, indiv, n = 3000, 1001, 3
    bs = (indiv + 3) // 4
    d_plink = torch.zeros(snps * bs, dtype=torch.uint8, device="cuda")
    assert gpu_lib.mb200_simulate_plink_device(d_plink.data_ptr(), snps, indiv, 5) == 0
    d2 = torch.zeros_like(d_plink)
    assert gpu_lib.mb200_simulate_plink_device(d2.data_ptr(), snps, indiv, 5) == 0
    assert torch.equal(d_plink, d2)
    plink = d_plink.cpu().numpy().reshape(snps, bs)
    codes = simulate.unpack_2bit(plink, bs * 4)
    assert not (codes == 1).any() and not codes[:, indiv:].any()
    Z = simulate.plink_decode(codes[:, :indiv]).T
    p = Z.mean(axis=0) / 2.0
    assert 0.08 < p.min() and p.max() < 0.62 and abs(p.mean() - 0.35) < 0.02
    h = C.c_void_p(None)
    assert gpu_lib.mb200_plink2compressed_device(d_plink.data_ptr(), snps, indiv, None, n,
                                                 C.byref(h)) == 0
    try:
        f = api.get_compressed_freq(h, snps)
        assert np.allclose(f, p, rtol=0, atol=1e-15)
        B = torch.randn(n, snps, dtype=torch.float64, device="cuda")  # = column-major snps x n
        Cd = torch.empty(n, indiv, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        assert gpu_lib.mb200_dgemm_compressed_device(0, h, n, B.data_ptr(), snps, Cd.data_ptr(),
                                                     indiv) == 0
        assert gpu_lib.mb200_sync(h) == 0
        want = cport.dgemm_compressed("N", plink, snps, indiv, f, B.cpu().numpy().T, True)
        assert relerr(Cd.cpu().numpy().T, want) <= TOL
    finally:
        api.free_compressed(h)


@pytest.mark.parametrize("chunks", [1, 2, 3, 5, 16])
@pytest.mark.parametrize("snps,indiv,ncol", [(5000, 700, 10), (1537, 300, 3), (20000, 2500, 10)])
def test_dgemm_host_pipeline_chunks(gpu_lib, snps, indiv, ncol, chunks):
    """The host-pointer call cuts the SNP axis into chunks to overlap PCIe with the kernels
    ('N': K ranges accumulate into C; 'T': row ranges): every chunking gives the oracle's result,
    with leading dimensions, centred, missings, fixed and adaptive digits."""
    from miraculix_b200 import api
    plink, Z, f = simulate.simulate(snps, indiv, seed=snps + chunks, missing=0.01)
    rng = np.random.default_rng(7)
    gpu_lib.mb200_set_pipeline(chunks, 0.0)
    h = api.init_compressed(plink, snps, indiv, f, ncol)
    try:
        for slices in (0, 7):
            gpu_lib.mb200_set_dgemm_mode(1, slices)
            for tr, k, m in (("N", snps, indiv), ("T", indiv, snps)):
                ldb, ldc = k + 3, m + 8
                Bbig = np.asfortranarray(rng.standard_normal((ldb, ncol)))
                Cbig = np.full((ldc, ncol), 7.25, order="F")
                gpu_lib.dgemm_compressed(tr.encode(), h, ncol, Bbig.ctypes.data, ldb,
                                         Cbig.ctypes.data, ldc)
                assert gpu_lib.mb200_last_status() == 0
                want = cport.dgemm_compressed(tr, plink, snps, indiv, f, Bbig[:k], True)
                assert relerr(Cbig[:m], want) <= TOL, (tr, slices)
                assert np.all(Cbig[m:] == 7.25)  # rows beyond m untouched
    finally:
        api.free_compressed(h)
        gpu_lib.mb200_set_pipeline(4, 1.0e7)
        gpu_lib.mb200_set_dgemm_mode(1, 0)


def test_dgemm_host_pipeline_precision_fallback(gpu_lib):
    """Chunked 'T' whose product cancels (centred Z^T 1 = 0): some chunk fails the 6-digit bound,
    the flag travels back with the result and the whole product is redone with 8 digits."""
    from miraculix_b200 import api
    snps, indiv, ncol = 16384, 20000, 3
    plink, Z, f = simulate.simulate(snps, indiv, seed=99)
    gpu_lib.mb200_set_pipeline(4, 0.0)
    gpu_lib.mb200_set_dgemm_mode(1, 0)
    h = api.init_compressed(plink, snps, indiv, f, ncol)
    sl = C.c_int(0)
    try:
        ones = np.ones((indiv, ncol), order="F")
        got = api.dgemm_compressed("T", h, ones, snps, indiv)
        assert gpu_lib.mb200_dgemm_last_precision(h, C.byref(sl)) == 2 and sl.value == 8
        want = cport.dgemm_compressed("T", plink, snps, indiv, f, ones, True)
        assert np.abs(got - want).max() <= TOL * float(Z.sum(axis=0).max())
        B = np.asfortranarray(np.random.default_rng(3).standard_normal((snps, ncol)))
        got = api.dgemm_compressed("N", h, B, snps, indiv)
        assert gpu_lib.mb200_dgemm_last_precision(h, C.byref(sl)) == 1 and sl.value == 6
        assert relerr(got, cport.dgemm_compressed("N", plink, snps, indiv, f, B, True)) <= TOL
    finally:
        api.free_compressed(h)
        gpu_lib.mb200_set_pipeline(4, 1.0e7)


@pytest.mark.parametrize("b_dev,c_dev", [(0, 0), (0, 1), (1, 0), (1, 1)])
def test_dgemm_mixed_host_device_panels(gpu_lib, b_dev, c_dev):
    """mb200_dgemm_compressed_mixed: each panel on the host or on the device, chunked pipeline
    forced on, both directions and both kernel families, against the oracle."""
    import torch
    from miraculix_b200 import api
    snps, indiv, ncol = 6000, 900, 5
    plink, Z, f = simulate.simulate(snps, indiv, seed=31, missing=0.01)
    rng = np.random.default_rng(32)
    gpu_lib.mb200_set_pipeline(3, 0.0)
    h = api.init_compressed(plink, snps, indiv, f, ncol)
    try:
        for mode in (1, 0):
            gpu_lib.mb200_set_dgemm_mode(mode, 0)
            for tr, k, m in ((0, snps, indiv), (1, indiv, snps)):
                B = np.asfortranarray(rng.standard_normal((k, ncol)))
                want = cport.dgemm_compressed("NT"[tr], plink, snps, indiv, f, B, True)
                tB = torch.from_numpy(np.ascontiguousarray(B.T))          # (ncol, k) = column-major k x ncol
                tC = torch.full((ncol, m), 3.5, dtype=torch.float64)
                if b_dev:
                    tB = tB.cuda()
                if c_dev:
                    tC = tC.cuda()
                rc = gpu_lib.mb200_dgemm_compressed_mixed(tr, h, ncol, C.c_void_p(tB.data_ptr()), k, b_dev,
                                                          C.c_void_p(tC.data_ptr()), m, c_dev)
                assert rc == 0, gpu_lib.mb200_last_message()
                gpu_lib.mb200_sync(h)
                got = tC.cpu().numpy().T
                assert relerr(got, want) <= TOL, (mode, tr)
    finally:
        api.free_compressed(h)
        gpu_lib.mb200_set_pipeline(4, 1.0e7)
        gpu_lib.mb200_set_dgemm_mode(1, 0)
