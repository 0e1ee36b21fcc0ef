"""GPU tests of the single-process multi-device handles behind the reference C ABI:
plink2compressed / dgemm_compressed / free_compressed / mb200_cg_solve / snp_multiply_gpu / mb200_grm
with MIRACULIX_NGPUS, CUDA_DEVICE=a,b or mb200_set_devices -- one process, plain ctypes calls, no
torch.distributed.

On a box with >= 2 GPUs the shards sit on different devices (peer-memory kernels over NVLink); on
a one-GPU box the device list names device 0 several times, which runs the same host threads,
events, exchange kernels and staging code with same-device "peers".

Bars: sharded results against the oracle <= 1e-10 (north_star), against the single-device result
<= 1e-12 (only the fp64 summation order of the partials differs); shard layouts, crossproduct
counts and GRMs bit-exact.
"""
import ctypes as C
import os

import numpy as np
import pytest

from oracle import cg as ocg
from oracle import cport, simulate

pytestmark = pytest.mark.gpu
TOL = 1e-10


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def device_count():
    import torch
    return torch.cuda.device_count()


def device_list(world):
    """world shards on distinct devices when the box has them, else cycling over what is there."""
    n = device_count()
    return [g % n for g in range(world)]


@pytest.fixture(autouse=True)
def _defaults(gpu_lib):
    from miraculix_b200 import api
    api.set_options(use_gpu=True, not_center=False)
    api.set_devices(None)
    yield
    api.set_devices(None)
    gpu_lib.mb200_set_dgemm_mode(1, 0)
    api.set_options(use_gpu=True, not_center=False)
    for k in ("MIRACULIX_NGPUS", "TILE_SIZE", "MB200_GRM_PASSES", "MB200_GRM_CHUNK_MB"):
        os.environ.pop(k, None)


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("snps,indiv,missing", [(2051, 777, 0.01), (1024, 203, 0.0), (5000, 1037, 0.02),
                                                (700, 64, 0.0)])
def test_shards_are_bit_exact_slices(gpu_lib, world, snps, indiv, missing):
    """Every shard is an ordinary handle of its SNP range: layouts and counts equal the oracle's
    repack of that slice of the bed; ranges are contiguous multiples of 512 and cover all SNPs."""
    from miraculix_b200 import api
    plink, Z, f = simulate.simulate(snps, indiv, seed=snps + world, missing=missing)
    api.set_devices(device_list(world))
    h = api.init_compressed(plink, snps, indiv, f, 2)
    try:
        sh = api.shards(h)
        assert 1 <= len(sh) <= world
        assert sh[0][1] == 0 and sum(c for _, _, c in sh) == snps
        pos = 0
        for sub, b, c in sh:
            assert b == pos and (b % 512 == 0) and c > 0
            pos += c
            sm, im, cnt = api.download_layout(sub)
            sm_o, im_o = cport.repack(plink[b:b + c], c, indiv)
            assert np.array_equal(sm, sm_o) and np.array_equal(im, im_o)
            assert np.array_equal(cnt.astype(np.int64), cport.allele_counts(plink[b:b + c], c, indiv))
        assert np.array_equal(api.get_compressed_freq(h, snps), f)
        if device_count() >= 2 and len(sh) >= 2:
            assert {api.get_layout(s).device for s, _, _ in sh} == set(device_list(len(sh)))
    finally:
        api.free_compressed(h)
    assert not h.value


MODES = [(1, 0), (1, 6), (1, 8), (0, 7)]


@pytest.mark.parametrize("mode,slices", MODES)
@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("snps,indiv,ncol", [(2051, 777, 3), (1003, 203, 10), (4097, 2050, 7), (40000, 9000, 4)])
def test_multi_dgemm_matches_oracle_and_single(gpu_lib, snps, indiv, ncol, world, mode, slices):
    """dgemm_compressed 'N' / 'T' on a sharded handle, plain pageable numpy panels."""
    from miraculix_b200 import api
    plink, Z, f = simulate.simulate(snps, indiv, seed=snps + ncol, missing=0.01)
    rng = np.random.default_rng(ncol + world)
    gpu_lib.mb200_set_dgemm_mode(mode, slices)
    h1 = api.init_compressed(plink, snps, indiv, f, ncol)
    api.set_devices(device_list(world))
    hm = api.init_compressed(plink, snps, indiv, f, ncol)
    try:
        assert gpu_lib.mb200_shard_count(hm) == min(world, (snps + 511) // 512)
        for tr, k in (("N", snps), ("T", indiv)):
            B = np.asfortranarray(rng.standard_normal((k, ncol)))
            got = api.dgemm_compressed(tr, hm, B, snps, indiv)
            one = api.dgemm_compressed(tr, h1, B, snps, indiv)
            assert relerr(got, one) <= 1e-12, (tr, mode)
            if snps * indiv <= 10_000_000:
                want = cport.dgemm_compressed(tr, plink, snps, indiv, f, B, True)
                assert relerr(got, want) <= TOL, (tr, mode)
    finally:
        api.free_compressed(hm)
        api.free_compressed(h1)


def test_multi_dgemm_leading_dimensions_not_centred_and_errors(gpu_lib):
    from miraculix_b200 import api
    snps, indiv, n = 3000, 500, 4
    plink, Z, f = simulate.simulate(snps, indiv, seed=3)
    rng = np.random.default_rng(4)
    api.set_options(use_gpu=True, not_center=True)
    api.set_devices(device_list(2))
    h = api.init_compressed(plink, snps, indiv, f, n)
    try:
        ldb, ldc = snps + 13, indiv + 5
        Bbig = np.asfortranarray(rng.standard_normal((ldb, n)))
        Cbig = np.full((ldc, n), 777.0, order="F")
        gpu_lib.dgemm_compressed(b"n", h, n, Bbig.ctypes.data, ldb, Cbig.ctypes.data, ldc)
        assert gpu_lib.mb200_last_status() == 0, gpu_lib.mb200_last_message()
        want = cport.dgemm_compressed("N", plink, snps, indiv, f, Bbig[:snps], False)
        assert relerr(Cbig[:indiv], want) <= TOL
        assert np.all(Cbig[indiv:] == 777.0)
        ldb, ldc = indiv + 3, snps + 7
        Bbig = np.asfortranarray(rng.standard_normal((ldb, n)))
        Cbig = np.full((ldc, n), -5.0, order="F")
        gpu_lib.dgemm_compressed(b"T", h, n, Bbig.ctypes.data, ldb, Cbig.ctypes.data, ldc)
        assert gpu_lib.mb200_last_status() == 0, gpu_lib.mb200_last_message()
        want = cport.dgemm_compressed("T", plink, snps, indiv, f, Bbig[:indiv], False)
        assert relerr(Cbig[:snps], want) <= TOL
        assert np.all(Cbig[snps:] == -5.0)
        # n > max_n is refused, C untouched; device-pointer entry points refuse a sharded handle
        B5 = np.asfortranarray(rng.standard_normal((snps, n + 1)))
        C5 = np.zeros((indiv, n + 1), order="F")
        gpu_lib.dgemm_compressed(b"N", h, n + 1, B5.ctypes.data, snps, C5.ctypes.data, indiv)
        assert gpu_lib.mb200_last_status() != 0 and not C5.any()
        assert gpu_lib.mb200_dgemm_compressed_device(0, h, 1, None, snps, None, indiv) != 0
    finally:
        api.free_compressed(h)


def test_multi_adaptive_check_uses_the_reduced_result(gpu_lib):
    """'N' partials that cancel across shards: B = (v, -v) pattern over the two SNP halves makes
    every shard's partial large and the reduced C small; the 6-digit bound must be judged on the
    REDUCED result and the product redone with 8 digits, still matching the oracle."""
    from miraculix_b200 import api
    snps, indiv, ncol = 32768, 12000, 3
    plink, Z, f = simulate.simulate(snps, indiv, seed=77)
    api.set_options(use_gpu=True, not_center=True)
    gpu_lib.mb200_set_dgemm_mode(1, 0)
    api.set_devices(device_list(2))
    h = api.init_compressed(plink, snps, indiv, f, ncol)
    sl = C.c_int(0)
    try:
        (s0, b0, c0), (s1, b1, c1) = api.shards(h)
        rng = np.random.default_rng(5)
        B = np.asfortranarray(rng.standard_normal((snps, ncol)))
        got = api.dgemm_compressed("N", h, B, snps, indiv)
        want = cport.dgemm_compressed("N", plink, snps, indiv, f, B, False)
        assert relerr(got, want) <= TOL
        assert gpu_lib.mb200_dgemm_last_precision(s0, C.byref(sl)) >= 0 and sl.value == 6
        # identical SNP halves with opposite weights: exact cancellation to ~0
        plink2 = np.concatenate([plink[:c0], plink[:c0]])
        f2 = np.concatenate([f[:c0], f[:c0]])
        api.free_compressed(h)
        h = api.init_compressed(plink2, 2 * c0, indiv, f2, ncol)
        v = rng.standard_normal((c0, ncol))
        B2 = np.asfortranarray(np.concatenate([v, -v * (1 + 1e-9)]))
        got = api.dgemm_compressed("N", h, B2, 2 * c0, indiv)
        sub = api.shards(h)[0][0]
        gpu_lib.mb200_dgemm_last_precision(sub, C.byref(sl))
        assert sl.value == 8
        want = cport.dgemm_compressed("N", plink2, 2 * c0, indiv, f2, B2, False)
        scale = np.abs(cport.dgemm_compressed("N", plink[:c0], c0, indiv, f[:c0], np.asfortranarray(v), False)).max()
        assert np.abs(got - want).max() <= TOL * scale
    finally:
        api.free_compressed(h)


def test_pageable_and_pinned_panels_agree(gpu_lib):
    """Pageable numpy panels (what Julia / Fortran pass) go through the pinned ring, page-locked
    panels directly: same results, single-device and sharded, several pipeline chunks."""
    import torch
    from miraculix_b200 import api
    snps, indiv, ncol = 70000, 3000, 6
    plink, Z, f = simulate.simulate(snps, indiv, seed=9)
    rng = np.random.default_rng(10)
    gpu_lib.mb200_set_pipeline(4, 0.0)
    try:
        for world in (1, 2):
            api.set_devices(device_list(world) if world > 1 else None)
            h = api.init_compressed(plink, snps, indiv, f, ncol)
            try:
                for tr, k, m in (("N", snps, indiv), ("T", indiv, snps)):
                    B = np.asfortranarray(rng.standard_normal((k, ncol)))
                    page = api.dgemm_compressed(tr, h, B, snps, indiv)
                    tB = torch.from_numpy(np.ascontiguousarray(B.T)).pin_memory()     # (ncol, k) = col-major k x ncol
                    tC = torch.zeros((ncol, m), dtype=torch.float64).pin_memory()
                    gpu_lib.dgemm_compressed(tr.encode(), h, ncol, tB.data_ptr(), k, tC.data_ptr(), m)
                    assert gpu_lib.mb200_last_status() == 0
                    assert relerr(tC.numpy().T, page) <= 1e-12, (world, tr)
                    assert relerr(page, cport.dgemm_compressed(tr, plink, snps, indiv, f, B, True)) <= TOL
            finally:
                api.free_compressed(h)
    finally:
        gpu_lib.mb200_set_pipeline(4, 1.0e7)


def test_env_ngpus_selects_the_sharded_path(gpu_lib):
    """MIRACULIX_NGPUS is what an unchanged Julia / Fortran caller sets (needs >= 2 GPUs)."""
    from miraculix_b200 import api
    if device_count() < 2:
        pytest.skip("needs two GPUs")
    snps, indiv = 3000, 300
    plink, Z, f = simulate.simulate(snps, indiv, seed=1)
    os.environ["MIRACULIX_NGPUS"] = "2"
    h = api.init_compressed(plink, snps, indiv, f, 1)
    try:
        assert gpu_lib.mb200_shard_count(h) == 2
        devs = [api.get_layout(s).device for s, _, _ in api.shards(h)]
        assert devs == [0, 1]
        B = np.asfortranarray(np.random.default_rng(0).standard_normal((snps, 1)))
        assert relerr(api.dgemm_compressed("N", h, B, snps, indiv),
                      cport.dgemm_compressed("N", plink, snps, indiv, f, B, True)) <= TOL
    finally:
        api.free_compressed(h)
        os.environ.pop("MIRACULIX_NGPUS", None)


def test_cuda_device_env_selects_the_device(gpu_lib):
    """CUDA_DEVICE=1 lands on device 1 (reference cuda_utils.cu:188-248; needs >= 2 GPUs), and a
    process can hold handles on two devices at once (per-device kernel attributes)."""
    from miraculix_b200 import api
    if device_count() < 2:
        pytest.skip("needs two GPUs")
    snps, indiv, ncol = 5000, 1500, 5
    plink, Z, f = simulate.simulate(snps, indiv, seed=2)
    B = np.asfortranarray(np.random.default_rng(1).standard_normal((snps, ncol)))
    want = cport.dgemm_compressed("N", plink, snps, indiv, f, B, True)
    old = os.environ.get("CUDA_DEVICE")
    handles = []
    try:
        for dev in (1, 0):
            os.environ["CUDA_DEVICE"] = str(dev)
            h = api.init_compressed(plink, snps, indiv, f, ncol)
            handles.append(h)
            assert api.get_layout(h).device == dev
        for h in handles:
            assert relerr(api.dgemm_compressed("N", h, B, snps, indiv), want) <= TOL
        M = api.snp_crossprod(simulate.pack_2bit(Z), snps, indiv, is_snpmajor=False)
        assert np.array_equal(M, (Z.astype(np.int64) @ Z.astype(np.int64).T).astype(np.float64))
    finally:
        for h in handles:
            api.free_compressed(h)
        if old is None:
            os.environ.pop("CUDA_DEVICE", None)
        else:
            os.environ["CUDA_DEVICE"] = old


def test_simulated_handle_does_not_depend_on_the_device_count(gpu_lib):
    from miraculix_b200 import api
    snps, indiv = 2600, 333
    h1 = api.simulate_compressed(snps, indiv, 99, 2)
    api.set_devices(device_list(3))
    hm = api.simulate_compressed(snps, indiv, 99, 2)
    try:
        sm1, im1, cnt1 = api.download_layout(h1)
        for sub, b, c in api.shards(hm):
            sm, im, cnt = api.download_layout(sub)
            assert np.array_equal(sm, sm1[b:b + c])
            assert np.array_equal(cnt, cnt1[b:b + c])
        assert np.array_equal(api.get_compressed_freq(hm, snps), api.get_compressed_freq(h1, snps))
        rng = np.random.default_rng(3)
        B = np.asfortranarray(rng.standard_normal((indiv, 2)))
        assert relerr(api.dgemm_compressed("T", hm, B, snps, indiv),
                      api.dgemm_compressed("T", h1, B, snps, indiv)) <= 1e-12
    finally:
        api.free_compressed(hm)
        api.free_compressed(h1)


@pytest.mark.parametrize("world", [2, 3])
def test_multi_cg_matches_single_and_oracle(gpu_lib, world):
    """mb200_cg_solve on a sharded handle: the reference script's loop on every device in
    lockstep, one peer-memory all-reduce per product."""
    from miraculix_b200 import api
    snps, indiv, iters = 3000, 400, 12
    plink, Z, f = simulate.simulate(snps, indiv, seed=31)
    rng = np.random.default_rng(32)
    b, x0 = rng.standard_normal(indiv), rng.standard_normal(indiv)
    lam = 0.05 * snps
    h1 = api.init_compressed(plink, snps, indiv, f, 1)
    api.set_devices(device_list(world))
    hm = api.init_compressed(plink, snps, indiv, f, 1)
    try:
        x1, it1, r1 = api.cg_solve(h1, b, x0, lam=lam, tol=0.0, max_iter=iters)
        xm, itm, rm = api.cg_solve(hm, b, x0, lam=lam, tol=0.0, max_iter=iters)
        assert it1 == itm == iters
        assert relerr(xm, x1) <= 1e-9
        xo, ito, reso = ocg.cg_solve(plink, snps, indiv, f, b, x0, lam=lam, tol=0.0, max_iter=iters)
        assert relerr(xm, xo) <= 1e-8
        # convergence-controlled run: same iteration count on every device (bit-identical scalars)
        tol = 1e-8 * np.linalg.norm(b)
        xc, itc, rc = api.cg_solve(hm, b, x0, lam=lam, tol=tol, max_iter=500)
        assert itc < 500 and rc < tol
        Gx = cport.dgemm_compressed("N", plink, snps, indiv, f,
                                    cport.dgemm_compressed("T", plink, snps, indiv, f, xc.reshape(-1, 1), True), True)
        assert np.linalg.norm(b - (Gx.ravel() + lam * xc)) <= 10 * tol
    finally:
        api.free_compressed(hm)
        api.free_compressed(h1)


def _grm_inputs(snps, indiv, seed):
    plink, Z, f = simulate.simulate(snps, indiv, seed=seed)
    return simulate.pack_2bit(Z), Z, f


@pytest.mark.parametrize("world", [1, 2, 3])
@pytest.mark.parametrize("snps,indiv", [(1000, 300), (4099, 1100), (777, 2051), (5000, 129)])
def test_tiled_crossproduct_counts_bit_exact(gpu_lib, world, snps, indiv):
    """snp_multiply_gpu through the tiled out-of-core path, several passes and chunks forced."""
    from miraculix_b200 import api
    x, Z, f = _grm_inputs(snps, indiv, snps + indiv)
    want = (Z.astype(np.int64) @ Z.astype(np.int64).T).astype(np.float64)
    api.set_devices(device_list(world) if world > 1 else None)
    for passes, tile, chunk in ((None, None, None), ("2", None, "0.5"), (None, "128", None)):
        for key, val in (("MB200_GRM_PASSES", passes), ("TILE_SIZE", tile), ("MB200_GRM_CHUNK_MB", chunk)):
            if val is None:
                os.environ.pop(key, None)
            else:
                os.environ[key] = val
        M = api.snp_crossprod(x, snps, indiv, is_snpmajor=False)
        assert np.array_equal(M, want), (world, passes, tile, chunk)


@pytest.mark.parametrize("world", [1, 2])
def test_tiled_grm_equals_device_resident_grm(gpu_lib, world):
    """mb200_grm (host pointers, tiled, any device count) against mb200_crossprod_device (fused
    epilogue, one device): the same exact integer expression -> bit-identical; both against the
    oracle's restatement of crossproduct.jl:91-106 to 1e-10."""
    import torch
    from miraculix_b200 import api
    snps, indiv = 6000, 1300
    x, Z, f = _grm_inputs(snps, indiv, 5)
    api.set_devices(device_list(world) if world > 1 else None)
    os.environ["MB200_GRM_PASSES"] = "2"
    G = api.grm(x, snps, indiv, do_scale=True, allele_freq=f)
    os.environ.pop("MB200_GRM_PASSES")
    api.set_devices(None)
    dx = torch.from_numpy(x).cuda()
    df = torch.from_numpy(f).cuda()
    dG = torch.empty(indiv * indiv, dtype=torch.float64, device="cuda")
    api._check(gpu_lib.mb200_crossprod_device(dx.data_ptr(), snps, indiv, 0, 1, df.data_ptr(), 1,
                                              dG.data_ptr()), "crossprod_device")
    torch.cuda.synchronize()
    assert np.array_equal(G, dG.cpu().numpy().reshape(indiv, indiv).T)
    M = (Z.astype(np.int64) @ Z.astype(np.int64).T).astype(np.float64)
    want = cport.grm_epilogue(M, f, True)
    assert relerr(G, want) <= TOL
    assert np.array_equal(G, G.T)


def test_nonfinite_b_gives_nan_columns(gpu_lib):
    """A NaN / Inf entry of B poisons its column of C (as the reference's fp64 FMA chain does:
    0 * NaN = NaN), the other columns stay correct; both kernel families, centred and not."""
    from miraculix_b200 import api
    snps, indiv, ncol = 2051, 777, 3
    plink, Z, f = simulate.simulate(snps, indiv, seed=8)
    rng = np.random.default_rng(9)
    for centred in (True, False):
        api.set_options(use_gpu=True, not_center=not centred)
        for mode in (1, 0):
            gpu_lib.mb200_set_dgemm_mode(mode, 0)
            h = api.init_compressed(plink, snps, indiv, f, ncol)
            try:
                for tr, k in (("N", snps), ("T", indiv)):
                    B = np.asfortranarray(rng.standard_normal((k, ncol)))
                    want = cport.dgemm_compressed(tr, plink, snps, indiv, f, B, centred)
                    for bad in (np.nan, np.inf):
                        Bb = B.copy(order="F")
                        Bb[k // 2, 1] = bad
                        got = api.dgemm_compressed(tr, h, Bb, snps, indiv)
                        assert not np.isfinite(got[:, 1]).any() if mode == 1 else np.isnan(got[:, 1]).any()
                        assert relerr(got[:, [0, 2]], want[:, [0, 2]]) <= TOL, (tr, mode, bad)
            finally:
                api.free_compressed(h)


def test_concurrent_handles_on_one_device(gpu_lib):
    """Four ordinary handles on one GPU driven from four host threads at once: the tensor-core
    kernels of different handles share the device and must not disturb each other.  (Regression:
    the TMA refill of a packed-Z slot could overtake the expanders' shared-memory loads of the
    previous chunk when other kernels changed the timing -- sporadic wrong rows; fixed by a proxy
    fence before the refill.)  Results must equal the sequential ones up to atomic summation order."""
    import threading
    from miraculix_b200 import api
    snps, indiv, ncol, nthreads = 125_000, 40_000, 10, 4
    hs = [api.simulate_compressed(snps, indiv, 500 + i, ncol) for i in range(nthreads)]
    rng = np.random.default_rng(7)
    B = {"N": np.asfortranarray(rng.standard_normal((snps, ncol))),
         "T": np.asfortranarray(rng.standard_normal((indiv, ncol)))}
    try:
        ref = [{d: api.dgemm_compressed(d, h, B[d], snps, indiv) for d in "NT"} for h in hs]
        for rep in range(6):
            out = [None] * nthreads

            def work(i):
                out[i] = {d: api.dgemm_compressed(d, hs[i], B[d], snps, indiv) for d in "NT"}
            th = [threading.Thread(target=work, args=(i,)) for i in range(nthreads)]
            [t.start() for t in th]
            [t.join() for t in th]
            for i in range(nthreads):
                for d in "NT":
                    assert relerr(out[i][d], ref[i][d]) <= 1e-12, (rep, i, d)
    finally:
        for h in hs:
            api.free_compressed(h)


@pytest.mark.parametrize("world", [2, 3])
def test_read_bed_and_device_bed_sharded(gpu_lib, tmp_path, world):
    """mb200_read_bed and mb200_plink2compressed_device with a device list: every device streams /
    copies its own SNP range; shards bit-exact, frequencies = read_bed's popcount rule (file) or the
    counts (device bed, no f given), products against the oracle."""
    import torch
    from oracle import bedfile
    from miraculix_b200 import api
    snps, indiv, ncol = 2600, 517, 3
    plink, Z, f = simulate.simulate(snps, indiv, seed=12, missing=0.02)
    path = bedfile.write_bed(str(tmp_path / "geno"), plink, snps, indiv)
    p_o, f_o, s_o, n_o = bedfile.read_bed(path)
    api.set_devices(device_list(world))
    h, fq, s, n = api.read_bed_compressed(path, ncol)
    d_plink = torch.from_numpy(plink).cuda()
    h2 = C.c_void_p(None)
    api._check(gpu_lib.mb200_plink2compressed_device(d_plink.data_ptr(), snps, indiv, None, ncol, C.byref(h2)),
               "plink2compressed_device")
    try:
        assert (s, n) == (snps, indiv) and np.array_equal(fq, f_o)
        for hh in (h, h2):
            assert gpu_lib.mb200_shard_count(hh) == world
            for sub, b, c in api.shards(hh):
                sm, im, cnt = api.download_layout(sub)
                sm_o, im_o = cport.repack(plink[b:b + c], c, indiv)
                assert np.array_equal(sm, sm_o) and np.array_equal(im, im_o)
        f2 = api.get_compressed_freq(h2, snps)
        assert np.array_equal(f2, cport.allele_counts(plink, snps, indiv) / (2.0 * indiv))
        B = np.asfortranarray(np.random.default_rng(1).standard_normal((snps, ncol)))
        assert relerr(api.dgemm_compressed("N", h, B, snps, indiv),
                      cport.dgemm_compressed("N", plink, snps, indiv, f_o, B, True)) <= TOL
        assert relerr(api.dgemm_compressed("N", h2, B, snps, indiv),
                      cport.dgemm_compressed("N", plink, snps, indiv, f2, B, True)) <= TOL
    finally:
        api.free_compressed(h)
        api.free_compressed(h2)
