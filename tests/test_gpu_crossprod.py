"""GPU parity tests for the crossproduct / GRM / LD path, through the C-ABI.

Bars: integer counts bit-exact; centred GRM and LD max rel. error <= 1e-10 against the oracle's
restatement of src/bindings/Julia/crossproduct.jl.  Shapes re-express
tests/crossproduct/test_grm.jl:114-157 (random 0/1/2 matrices, "uneven" dimensions
snps in {953, 10251} x indiv in {752, 5343}).
"""
import os

import numpy as np
import pytest

from oracle import cport, simulate

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
TOL = 1e-10


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def exact_counts(vals):
    """Z Z^T via fp64 BLAS: exact while entries stay below 2^53 (test_grm.jl uses BLAS.gemm)."""
    v = vals.astype(np.float64)
    return v @ v.T


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("snps,indiv", [(953, 752), (64, 5), (129, 257), (2051, 1343),
                                        (10251, 5343), (1024, 515), (8176, 1030)])
def test_counts_bit_exact(gpu_lib, snps, indiv, mode):
    from miraculix_b200 import api
    if mode == 0 and snps * indiv > 4e6:
        pytest.skip("SIMT cross-check only on the small shapes")
    _, Z, _ = simulate.simulate(snps, indiv, seed=snps)
    x = simulate.pack_2bit(Z)
    gpu_lib.mb200_set_crossprod_mode(mode)
    try:
        M = api.snp_crossprod(x, snps, indiv, is_snpmajor=False)
    finally:
        gpu_lib.mb200_set_crossprod_mode(2)
    assert np.array_equal(M, M.T)  # crossproduct.jl:61
    if snps * indiv <= 4e6:
        assert np.array_equal(M, cport.crossprod(x, snps, indiv))
    assert np.array_equal(M, exact_counts(Z))


@pytest.mark.parametrize("mode", [0, 1, 2])
def test_plink_format_table_and_missing_quirk(gpu_lib, mode):
    """is_plink_format=True: bytes go through the reference table; a byte with a missing code
    counts as four 3s, exactly like the reference."""
    from miraculix_b200 import api
    snps, indiv = 777, 130
    plink, Z, _ = simulate.simulate(snps, indiv, seed=2, missing=0.03)
    codes = simulate.unpack_2bit(plink, indiv).T
    x = simulate.pack_2bit(codes)  # individual-major, PLINK-coded
    gpu_lib.mb200_set_crossprod_mode(mode)
    try:
        M = api.snp_crossprod(x, snps, indiv, is_snpmajor=False, is_plink_format=True)
    finally:
        gpu_lib.mb200_set_crossprod_mode(2)
    assert np.array_equal(M, cport.crossprod(x, snps, indiv, True))
    lut = np.load(os.path.join(GOLDEN, "plink2twobit_lut.npy"))
    vals = simulate.unpack_2bit(lut[x], x.shape[1] * 4)
    assert np.array_equal(M, exact_counts(vals))


def test_k_slabs_accumulate_exactly(gpu_lib, monkeypatch):
    """Force several K-slabs (int32 += epilogue) and compare with the single-slab result."""
    from miraculix_b200 import api
    snps, indiv = 5000, 600
    _, Z, _ = simulate.simulate(snps, indiv, seed=4)
    x = simulate.pack_2bit(Z)
    monkeypatch.setenv("MB200_SLAB_GB", "0.0005")  # ~0.5 MB of int8 per slab -> many slabs
    M = api.snp_crossprod(x, snps, indiv, is_snpmajor=False)
    assert np.array_equal(M, exact_counts(Z))


@pytest.mark.parametrize("snps,indiv", [(953, 752), (3001, 1025)])
def test_grm_matches_oracle(gpu_lib, snps, indiv):
    from miraculix_b200 import api
    _, Z, f = simulate.simulate(snps, indiv, seed=indiv)
    x = simulate.pack_2bit(Z)
    counts = cport.crossprod(x, snps, indiv)
    for do_scale in (True, False):
        G = api.grm(x, snps, indiv, do_scale=do_scale, allele_freq=f)
        assert relerr(G, cport.grm_epilogue(counts, f, do_scale)) <= TOL
        assert relerr(G, G.T) <= 1e-14
    # test_grm.jl:133-139: equals the centred dense product
    Zc = Z - 2.0 * f[None, :]
    G = api.grm(x, snps, indiv, do_scale=True, allele_freq=f)
    assert relerr(G, Zc @ Zc.T / (2.0 * np.sum(f * (1.0 - f)))) <= 1e-9


@pytest.mark.parametrize("snps,indiv", [(2051, 1343), (9300, 3000)])
def test_grm_identical_over_gemm_paths(gpu_lib, snps, indiv):
    """SIMT (0), one CTA per tile (1) and CTA pairs (2, tcgen05 cta_group::2) produce the same counts
    and -- the centring is exact integer arithmetic with one final scale -- bit-identical, exactly
    symmetric GRMs."""
    from miraculix_b200 import api
    _, Z, f = simulate.simulate(snps, indiv, seed=snps + 1)
    x = simulate.pack_2bit(Z)
    res = []
    try:
        for mode in (0, 1, 2):
            gpu_lib.mb200_set_crossprod_mode(mode)
            res.append((api.snp_crossprod(x, snps, indiv, is_snpmajor=False),
                        api.grm(x, snps, indiv, do_scale=True, allele_freq=f)))
    finally:
        gpu_lib.mb200_set_crossprod_mode(2)
    for M, G in res[1:]:
        assert np.array_equal(M, res[0][0])
        assert np.array_equal(G, res[0][1])
    assert np.array_equal(res[2][1], res[2][1].T)
    want = cport.grm_epilogue(exact_counts(Z), f, True)
    assert relerr(res[2][1], want) <= TOL


def test_ld_matches_oracle(gpu_lib):
    from miraculix_b200 import api
    snps, indiv = 600, 2000
    plink, Z, f = simulate.simulate(snps, indiv, seed=6)
    x = simulate.pack_2bit(Z.T)  # SNP-major twobit
    R = api.ld(x, snps, indiv, allele_freq=f)
    counts = cport.crossprod(x, indiv, snps)
    assert relerr(R, cport.ld_epilogue(counts, indiv, f)) <= TOL
    assert relerr(R, np.corrcoef(Z.T.astype(np.float64))) <= 1e-9


def test_grm_properties_full_size(gpu_lib):
    """BASELINE config 2 (50k SNPs x 10k individuals): size-independent checks --
    symmetry, diagonal = sum z^2, sampled entries against direct dot products, and the row sums
    M 1 = Z (Z^T 1) computed independently by dgemm_compressed."""
    import ctypes as C
    import torch
    from miraculix_b200 import api
    snps, indiv = 50000, 10000
    bs = (indiv + 3) // 4
    d_plink = torch.zeros(snps * bs, dtype=torch.uint8, device="cuda")
    assert gpu_lib.mb200_simulate_plink_device(d_plink.data_ptr(), snps, indiv, 2) == 0
    h = C.c_void_p(None)
    assert gpu_lib.mb200_plink2compressed_device(d_plink.data_ptr(), snps, indiv, None, 1,
                                                 C.byref(h)) == 0
    try:
        lay = api.get_layout(h)
        d_ans = torch.empty(indiv * indiv, dtype=torch.float64, device="cuda")
        # the handle's individual-major layout is value-coded 2-bit with a 128-byte pitch: view it
        # as a (indiv x 4*pitch) twobit matrix (padding fields are zero)
        rc = gpu_lib.mb200_crossprod_device(lay.d_indiv_major, lay.pitch_indiv_major * 4, indiv, 0,
                                            0, None, 0, d_ans.data_ptr())
        assert rc == 0
        M = d_ans.view(indiv, indiv)
        assert torch.equal(M, M.T)
        _, im, cnt = api.download_layout(h)
        rng = np.random.default_rng(0)
        rows = rng.integers(0, indiv, size=24)
        Zs = cport.unpack_device(im[rows], snps).astype(np.int64)
        sub = M[torch.as_tensor(rows, device="cuda")][:, torch.as_tensor(rows, device="cuda")]
        assert np.array_equal(sub.cpu().numpy(), (Zs @ Zs.T).astype(np.float64))
        # row sums through the other kernel family (uncentred dgemm_compressed)
        api.set_options(use_gpu=True, not_center=True)
        ones = np.ones((indiv, 1))
        zt1 = api.dgemm_compressed("T", h, ones, snps, indiv)
        assert np.array_equal(zt1[:, 0], cnt.astype(np.float64))
        m = api.dgemm_compressed("N", h, zt1, snps, indiv)[:, 0]
        assert np.array_equal(M.sum(dim=0).cpu().numpy(), m)
    finally:
        api.set_options(use_gpu=True, not_center=False)
        api.free_compressed(h)


@pytest.mark.parametrize("snps,indiv,nparts", [(1003, 700, 3), (953, 1500, 8), (300, 100, 2)])
def test_grm_partition_assembles_to_full(gpu_lib, snps, indiv, nparts):
    """2-D output-tile partition: the panels of all partitions, mirrored, equal the
    single-GPU result bit for bit -- raw counts and the centred/scaled GRM (whose row sums the
    partitions obtain exactly from the input, without exchanging anything)."""
    from miraculix_b200 import api
    plink, Z, f = simulate.simulate(snps, indiv, seed=snps + nparts)
    x = simulate.pack_2bit(Z)
    for mode in (0, 1):
        full = api.snp_crossprod(x, snps, indiv, is_snpmajor=False) if mode == 0 else \
            api.grm(x, snps, indiv, do_scale=True, allele_freq=f)
        got = np.full((indiv, indiv), np.nan)
        seen = 0
        for part in range(nparts):
            rows, panel = api.grm_partition(x, snps, indiv, part, nparts, mode=mode, do_scale=True,
                                            allele_freq=f)
            for lr, i in enumerate(rows):
                if i >= indiv:
                    assert not panel[lr].any()  # padding rows of the last tile
                    continue
                first = (i // 256) * 256
                assert not panel[lr, :first].any()
                got[i, first:] = panel[lr, first:]
                seen += 1
        assert seen == indiv  # every row owned exactly once
        iu = np.triu_indices(indiv)
        assert np.array_equal(got[iu], full[iu]), mode


@pytest.mark.parametrize("snps,indiv,nparts", [(1003, 700, 3), (515, 1300, 2)])
def test_grm_partition_host_assembly(gpu_lib, snps, indiv, nparts):
    """sharded.assemble_grm_panels over the device panels == the single-GPU GRM, bit for bit, and
    == the oracle within the fp64 tolerance of north_star (1e-10 relative)."""
    from miraculix_b200 import api
    from miraculix_b200.sharded import assemble_grm_panels
    plink, Z, f = simulate.simulate(snps, indiv, seed=7 * snps + nparts)
    x = simulate.pack_2bit(Z)
    full = api.grm(x, snps, indiv, do_scale=True, allele_freq=f)
    panels = [api.grm_partition(x, snps, indiv, p, nparts, mode=1, do_scale=True, allele_freq=f)[1]
              for p in range(nparts)]
    got = assemble_grm_panels(panels, indiv)
    iu = np.triu_indices(indiv)
    assert np.array_equal(got[iu], full[iu])
    assert np.array_equal(got, got.T)
    assert relerr(got, cport.grm_epilogue(cport.crossprod(x, snps, indiv), f, True)) <= TOL


def test_grm_partition_with_k_slabs(gpu_lib, monkeypatch):
    """The configs[4] schedule (1M SNPs: the int8 expansion does not fit, K goes in slabs with int32
    partials per partition, the fused epilogue runs on the last slab) at test size: panels of a
    slabbed run are bit-identical to the single-slab panels, counts and centred GRM."""
    from miraculix_b200 import api
    snps, indiv, nparts = 4100, 900, 3
    _, Z, f = simulate.simulate(snps, indiv, seed=91)
    x = simulate.pack_2bit(Z)
    for mode in (0, 1):
        one = [api.grm_partition(x, snps, indiv, p, nparts, mode=mode, do_scale=True, allele_freq=f)
               for p in range(nparts)]
        monkeypatch.setenv("MB200_SLAB_GB", "0.0005")
        try:
            many = [api.grm_partition(x, snps, indiv, p, nparts, mode=mode, do_scale=True,
                                      allele_freq=f) for p in range(nparts)]
        finally:
            monkeypatch.delenv("MB200_SLAB_GB")
        for (r1, p1), (r2, p2) in zip(one, many):
            assert np.array_equal(r1, r2)
            assert np.array_equal(p1, p2), mode
    # and against the oracle through the first partition's first row tile
    rows, panel = many[0]
    want = cport.grm_epilogue(exact_counts(Z), f, True)
    assert relerr(panel[:128, :], want[rows[:128], :] * (np.arange(indiv)[None, :] >= 0)) <= TOL


@pytest.mark.parametrize("snps,indiv,nparts", [(517, 1300, 2), (2100, 2900, 4), (300, 100, 2),
                                               (9300, 3000, 2), (1200, 5100, 8)])
def test_grm_partition_pairs_match_single_cta(gpu_lib, snps, indiv, nparts, monkeypatch):
    """CTA-pair kernel on partition panels: a pair takes two consecutive row tiles of the partition
    (mi, mi + nparts), so its second half can lie left of the diagonal (dropped) or past the last
    row tile (odd tile count).  Panels must equal the one-CTA-per-tile panels bit for bit, counts
    and centred GRM, single slab (with the K-split tail) and K slabs."""
    from miraculix_b200 import api
    _, Z, f = simulate.simulate(snps, indiv, seed=snps + indiv)
    x = simulate.pack_2bit(Z)
    for slab in (None, "0.002"):
        if slab:
            monkeypatch.setenv("MB200_SLAB_GB", slab)
        try:
            for mode in (0, 1):
                for part in range(nparts):
                    gpu_lib.mb200_set_crossprod_mode(1)
                    try:
                        r1, p1 = api.grm_partition(x, snps, indiv, part, nparts, mode=mode,
                                                   do_scale=True, allele_freq=f)
                    finally:
                        gpu_lib.mb200_set_crossprod_mode(2)
                    r2, p2 = api.grm_partition(x, snps, indiv, part, nparts, mode=mode, do_scale=True,
                                               allele_freq=f)
                    assert np.array_equal(r1, r2)
                    assert np.array_equal(p1, p2), (mode, part, slab)
        finally:
            if slab:
                monkeypatch.delenv("MB200_SLAB_GB")
    if True:  # counts of the last partition against the exact integer product
        rows, panel = api.grm_partition(x, snps, indiv, nparts - 1, nparts, mode=0)
        want = exact_counts(Z)
        for lr, i in enumerate(rows):
            if i < indiv:
                first = (i // 256) * 256
                assert np.array_equal(panel[lr, first:], want[i, first:])
                assert not panel[lr, :first].any()


def test_k_split_tail_matches_unsplit(gpu_lib, monkeypatch):
    """Shape whose upper-triangle tile count (156) leaves a partial last round on 148 SMs: those
    tiles are cut along K into pieces (int32 partials in a scratch tile + tail_finish_kernel).
    Counts must stay exact and the centred GRM bit-identical to the unsplit schedule; the
    partition panels go through the same path."""
    from miraculix_b200 import api
    snps, indiv = 9300, 3000
    _, Z, f = simulate.simulate(snps, indiv, seed=77)
    x = simulate.pack_2bit(Z)
    M = api.snp_crossprod(x, snps, indiv, is_snpmajor=False)
    assert np.array_equal(M, exact_counts(Z))
    G = api.grm(x, snps, indiv, do_scale=True, allele_freq=f)
    rows, panel = api.grm_partition(x, snps, indiv, 1, 2, mode=1, do_scale=True, allele_freq=f)
    monkeypatch.setenv("MB200_NO_TAIL_SPLIT", "1")
    G0 = api.grm(x, snps, indiv, do_scale=True, allele_freq=f)
    rows0, panel0 = api.grm_partition(x, snps, indiv, 1, 2, mode=1, do_scale=True, allele_freq=f)
    assert np.array_equal(G, G0)
    assert np.array_equal(panel, panel0)
    want = cport.grm_epilogue(exact_counts(Z), f, True)
    assert relerr(G, want) <= TOL


def _device_rows(ptr, pitch, rows):
    """Rows of a library-owned packed device matrix as a (len(rows), pitch) uint8 numpy array."""
    import torch

    class _H:
        pass
    out = np.empty((len(rows), pitch), dtype=np.uint8)
    for t, r in enumerate(rows):
        hold = _H()
        hold.__cuda_array_interface__ = {"shape": (pitch,), "typestr": "|u1", "data": (int(ptr) + int(r) * pitch, False),
                                         "version": 3, "strides": None}
        out[t] = torch.as_tensor(hold, device="cuda").cpu().numpy()
    return out


def test_grm_partition_full_size_sampled_parity(gpu_lib):
    """BASELINE configs[4] at FULL size: one partition (3 of 8) of the centred+scaled GRM of 1M SNPs x
    100k individuals -- K slabs, int32 partial panel, CTA pairs with dead halves, all at once.  32
    sampled panel entries against direct int64 dot products of the decoded packed rows and the update
    of crossproduct.jl:91-106 in long double, with the row sums m = Z (Z^T 1) obtained independently
    through dgemm_compressed (not centred; integer-exact).  Skips without ~120 GB of free HBM."""
    import ctypes as C
    import torch
    from miraculix_b200 import api
    snps, indiv, part, nparts = 1_000_000, 100_000, 3, 8
    free_b, _ = torch.cuda.mem_get_info()
    if free_b < 120e9:
        pytest.skip("needs ~120 GB of free device memory")
    h = api.simulate_compressed(snps, indiv, 20261105, 1)
    try:
        lay = api.get_layout(h)
        pitch = lay.pitch_indiv_major
        kcols = pitch * 4
        f = api.get_compressed_freq(h, snps)
        d_f = torch.zeros(kcols, dtype=torch.float64, device="cuda")
        d_f[:snps] = torch.from_numpy(f).cuda()
        lrows = gpu_lib.mb200_crossprod_partition_rows(indiv, part, nparts)
        panel = torch.empty(lrows * indiv, dtype=torch.float64, device="cuda")
        api._check(gpu_lib.mb200_crossprod_partition_device(lay.d_indiv_major, kcols, indiv, 0, 1, d_f.data_ptr(), 1,
                                                            part, nparts, panel.data_ptr()), "partition")
        torch.cuda.synchronize()
        P = panel.view(lrows, indiv)
        # exact row sums through the other kernel family
        api.set_options(use_gpu=True, not_center=True)
        colsum = api.dgemm_compressed("T", h, np.ones((indiv, 1)), snps, indiv)
        m = api.dgemm_compressed("N", h, colsum, snps, indiv)[:, 0]
        total, c = np.longdouble(m.sum()), np.longdouble(2.0 * (f * (1.0 - f)).sum())
        rng = np.random.default_rng(4)
        lr = rng.integers(0, lrows, size=32)
        gi = ((lr // 128) * nparts + part) * 128 + lr % 128
        lr, gi = lr[gi < indiv], gi[gi < indiv]
        first = gi // 256 * 256
        jj = first + (rng.random(gi.size) * (indiv - first)).astype(np.int64)
        jj[0] = gi[0]                                   # a diagonal entry
        jj[1] = indiv - 1
        shifts = np.arange(4, dtype=np.uint8) * 2
        Zi = _device_rows(lay.d_indiv_major, pitch, gi)
        Zj = _device_rows(lay.d_indiv_major, pitch, jj)
        got = P[torch.as_tensor(lr, device="cuda"), torch.as_tensor(jj, device="cuda")].cpu().numpy()
        worst = 0.0
        for t in range(gi.size):
            zi = ((Zi[t][:, None] >> shifts) & 3).reshape(-1).astype(np.int64)
            zj = ((Zj[t][:, None] >> shifts) & 3).reshape(-1).astype(np.int64)
            cnt = np.longdouble(int(zi @ zj))
            want = (cnt - (np.longdouble(m[gi[t]]) + np.longdouble(m[jj[t]])) / indiv + total / indiv / indiv) / c
            worst = max(worst, abs(float(got[t] - want)))
        assert worst <= TOL * float(np.abs(got).max()), worst
    finally:
        api.set_options(use_gpu=True, not_center=False)
        api.free_compressed(h)


def test_snp_multiply_out_of_core_large(gpu_lib):
    """snp_multiply_gpu with more individuals than the in-core path could hold next to its int8 slab
    and int32 matrix in one piece: 120k individuals (115 GB of fp64 result, in HOST memory only), device
    memory bounded by TILE_SIZE.  Sampled entries, both triangles, against int64 dot products.  Skips
    without ~150 GB of free host memory."""
    import ctypes as C
    import os
    from miraculix_b200 import api
    indiv, snps = 120_000, 4096
    need = 8 * indiv * indiv
    avail = 0
    for ln in open("/proc/meminfo"):
        if ln.startswith("MemAvailable:"):
            avail = int(ln.split()[1]) * 1024
    if avail < 1.3 * need:
        pytest.skip("needs ~150 GB of free host memory")
    rng = np.random.default_rng(6)
    x = rng.integers(0, 256, size=(indiv, snps // 4), dtype=np.uint8)
    x &= rng.integers(0, 256, size=(indiv, snps // 4), dtype=np.uint8)   # thin out: more zeros
    x[(x & 0x03) == 0x03] &= 0xFE                                         # field value 3 -> 2 (values are 0..2)
    x[(x & 0x0C) == 0x0C] &= 0xFB
    x[(x & 0x30) == 0x30] &= 0xEF
    x[(x & 0xC0) == 0xC0] &= 0xBF
    M = np.empty((indiv, indiv))
    os.environ["TILE_SIZE"] = "8192"
    try:
        rc = gpu_lib.snp_multiply_gpu(x.ctypes.data, snps, indiv, M.ctypes.data, False)
    finally:
        os.environ.pop("TILE_SIZE", None)
    assert rc == 0, gpu_lib.mb200_last_message()
    shifts = np.arange(4, dtype=np.uint8) * 2
    ii = rng.integers(0, indiv, size=48)
    jj = rng.integers(0, indiv, size=48)
    ii[:3], jj[:3] = (0, indiv - 1, 777), (indiv - 1, indiv - 1, 777)
    for i, j in zip(ii.tolist(), jj.tolist()):
        zi = ((x[i][:, None] >> shifts) & 3).reshape(-1).astype(np.int64)
        zj = ((x[j][:, None] >> shifts) & 3).reshape(-1).astype(np.int64)
        want = float(zi @ zj)
        assert M[i, j] == want and M[j, i] == want, (i, j)
