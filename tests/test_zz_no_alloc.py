"""GPU test of mb200_dgemm_compressed_noalloc (include/miraculix_b200.h): the fp64 dgemm_compressed
on device pointers whose scratch (the centring scalars) sits in a static __device__ array, so the
call itself allocates nothing.  No counterpart in the reference;
checked through the C ABI against the existing entry point
mb200_dgemm_compressed_device and the oracle, at the tolerance of the other dgemm tests (1e-10),
on two calls in a row that share the scratch.

The file name sorts last on purpose.
"""
import ctypes as C

import numpy as np
import pytest

from oracle import cport, simulate

pytestmark = pytest.mark.gpu
TOL = 1e-10


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


@pytest.fixture(autouse=True)
def _defaults(gpu_lib):
    from miraculix_b200 import api
    api.set_options(use_gpu=True, not_center=False)
    api.set_devices(None)
    gpu_lib.mb200_set_dgemm_mode(1, 0)
    yield
    api.set_devices(None)
    gpu_lib.mb200_set_dgemm_mode(1, 0)
    api.set_options(use_gpu=True, not_center=False)


def _colmajor(t):
    """torch (n, k) row-major == column-major k x n"""
    return t.cpu().numpy().T


@pytest.mark.parametrize("not_center", [False, True])
def test_noalloc_matches_existing_entry_point_twice_in_a_row(gpu_lib, not_center):
    import torch
    from miraculix_b200 import api
    snps, indiv, n1, n2 = 3000, 1001, 12, 5
    plink, Z, f = simulate.simulate(snps, indiv, seed=11, missing=0.01)
    api.set_options(use_gpu=True, not_center=not_center)
    h = api.init_compressed(plink, snps, indiv, f, n1)
    try:
        g = torch.Generator(device="cuda").manual_seed(3)
        pad = 7  # leading dimensions larger than the panels; the padding rows must stay untouched
        Bn = [torch.randn(n, snps + pad, dtype=torch.float64, device="cuda", generator=g) * s
              for n, s in ((n1, 1000.0), (n2, 1.0))]   # call 1 leaves LARGE scalars in the scratch
        Bt = [torch.randn(n, indiv + pad, dtype=torch.float64, device="cuda", generator=g) * s
              for n, s in ((n1, 1000.0), (n2, 1.0))]
        Cn = [torch.full((n, indiv + pad), 777.0, dtype=torch.float64, device="cuda") for n in (n1, n2)]
        Ct = [torch.full((n, snps + pad), 777.0, dtype=torch.float64, device="cuda") for n in (n1, n2)]
        Rn = [torch.empty_like(c) for c in Cn]
        Rt = [torch.empty_like(c) for c in Ct]
        torch.cuda.synchronize()

        def two_calls_each():
            for i, n in enumerate((n1, n2)):  # 'N' twice in a row, then 'T' twice in a row; no sync between
                rc = gpu_lib.mb200_dgemm_compressed_noalloc(0, h, n, Bn[i].data_ptr(), snps + pad,
                                                            Cn[i].data_ptr(), indiv + pad)
                assert rc == 0, gpu_lib.mb200_last_message()
            for i, n in enumerate((n1, n2)):
                rc = gpu_lib.mb200_dgemm_compressed_noalloc(1, h, n, Bt[i].data_ptr(), indiv + pad,
                                                            Ct[i].data_ptr(), snps + pad)
                assert rc == 0, gpu_lib.mb200_last_message()
            assert gpu_lib.mb200_sync(h) == 0

        two_calls_each()                      # first use: symbol address, kernels loaded
        free0 = torch.cuda.mem_get_info()[0]
        two_calls_each()                      # the checked calls
        assert torch.cuda.mem_get_info()[0] == free0, "the no-alloc entry point changed free device memory"

        # the existing entry point on the same inputs: fp64 kernel (mode 0) and the default tensor-core path
        for mode in (0, 1):
            gpu_lib.mb200_set_dgemm_mode(mode, 0)
            for i, n in enumerate((n1, n2)):
                Rn[i].fill_(777.0)
                Rt[i].fill_(777.0)
                torch.cuda.synchronize()
                assert gpu_lib.mb200_dgemm_compressed_device(0, h, n, Bn[i].data_ptr(), snps + pad,
                                                             Rn[i].data_ptr(), indiv + pad) == 0
                assert gpu_lib.mb200_dgemm_compressed_device(1, h, n, Bt[i].data_ptr(), indiv + pad,
                                                             Rt[i].data_ptr(), snps + pad) == 0
                assert gpu_lib.mb200_sync(h) == 0
                assert relerr(_colmajor(Cn[i])[:indiv], _colmajor(Rn[i])[:indiv]) <= TOL, (mode, i)
                assert relerr(_colmajor(Ct[i])[:snps], _colmajor(Rt[i])[:snps]) <= TOL, (mode, i)
        # and the oracle
        for i in range(2):
            want = cport.dgemm_compressed("N", plink, snps, indiv, f, _colmajor(Bn[i])[:snps], not not_center)
            assert relerr(_colmajor(Cn[i])[:indiv], want) <= TOL
            want = cport.dgemm_compressed("T", plink, snps, indiv, f, _colmajor(Bt[i])[:indiv], not not_center)
            assert relerr(_colmajor(Ct[i])[:snps], want) <= TOL
            assert np.all(_colmajor(Cn[i])[indiv:] == 777.0) and np.all(_colmajor(Ct[i])[snps:] == 777.0)
    finally:
        api.free_compressed(h)


def test_noalloc_limits_and_errors(gpu_lib):
    import torch
    from miraculix_b200 import api
    snps, indiv = 1024, 203
    plink, Z, f = simulate.simulate(snps, indiv, seed=5)
    limit = gpu_lib.mb200_noalloc_max_cols()
    assert limit == 1024
    h = api.init_compressed(plink, snps, indiv, f, 2)
    try:
        # n is bound by the static scratch, not by the handle's max_n (= 2): 40 columns work
        n = 40
        B = torch.randn(n, snps, dtype=torch.float64, device="cuda")
        Cd = torch.zeros(n, indiv, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        assert gpu_lib.mb200_dgemm_compressed_noalloc(0, h, n, B.data_ptr(), snps, Cd.data_ptr(), indiv) == 0
        assert gpu_lib.mb200_sync(h) == 0
        want = cport.dgemm_compressed("N", plink, snps, indiv, f, _colmajor(B), True)
        assert relerr(_colmajor(Cd), want) <= TOL
        # beyond the limit: refused with a clear message, C untouched
        Cd.fill_(-3.0)
        torch.cuda.synchronize()
        rc = gpu_lib.mb200_dgemm_compressed_noalloc(0, h, limit + 1, B.data_ptr(), snps, Cd.data_ptr(), indiv)
        assert rc != 0 and gpu_lib.mb200_last_status() == rc
        msg = gpu_lib.mb200_last_message().decode()
        assert str(limit) in msg and "scratch" in msg, msg
        torch.cuda.synchronize()
        assert bool((Cd == -3.0).all())
        # bad leading dimension, null panels, invalid handle
        assert gpu_lib.mb200_dgemm_compressed_noalloc(0, h, 1, B.data_ptr(), snps - 1, Cd.data_ptr(), indiv) != 0
        assert gpu_lib.mb200_dgemm_compressed_noalloc(1, h, 1, None, indiv, None, snps) != 0
        assert gpu_lib.mb200_dgemm_compressed_noalloc(0, None, 1, B.data_ptr(), snps, Cd.data_ptr(), indiv) != 0
        assert gpu_lib.mb200_dgemm_compressed_noalloc(0, h, 0, None, snps, None, indiv) == 0
    finally:
        api.free_compressed(h)
    # a handle sharded over several devices is refused (device pointers belong to one device)
    api.set_devices([0, 0])
    hm = api.init_compressed(plink, snps, indiv, f, 2)
    try:
        assert gpu_lib.mb200_dgemm_compressed_noalloc(0, hm, 1, B.data_ptr(), snps, Cd.data_ptr(), indiv) != 0
    finally:
        api.free_compressed(hm)
        api.set_devices(None)
