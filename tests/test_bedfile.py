"""PLINK .bed ingest: the validation half runs on the CPU (no device is
touched before the file has been checked), the streaming + GPU repack half is marked gpu.
Oracle: oracle/bedfile.py, the numpy restatement of read_bed (read_plink.jl:161-228)."""
import numpy as np
import pytest

from oracle import bedfile, cport, simulate


def _triple(tmp_path, snps, indiv, seed=0, missing=0.0, **kw):
    plink, Z, f = simulate.simulate(snps, indiv, seed=seed, missing=missing)
    return bedfile.write_bed(str(tmp_path / "geno"), plink, snps, indiv, **kw), plink, Z


def test_bed_shape_and_oracle_reader(lib, tmp_path):
    from miraculix_b200 import api
    path, plink, Z = _triple(tmp_path, 37, 13, seed=1, missing=0.1)
    assert api.bed_shape(path) == (37, 13)
    p2, f2, s, n = bedfile.read_bed(path)
    assert (s, n) == (37, 13) and np.array_equal(p2, plink)
    # set-bit rule: 10 -> 1, 11 -> 2 and a missing 01 also counts one bit (read_plink.jl:197-203)
    codes = np.stack([(plink >> (2 * t)) & 3 for t in range(4)], axis=2).reshape(37, -1)[:, :13]
    bits = (codes & 1) + (codes >> 1)
    assert np.allclose(f2, bits.sum(axis=1) / 26.0, rtol=0, atol=0)


@pytest.mark.parametrize("case,code", [("suffix", -20), ("nofam", -21), ("magic", -23),
                                       ("small", -25), ("large", -26)])
def test_bed_errors_follow_read_bed(lib, tmp_path, case, code):
    """Same refusals as read_bed's error()/assert calls, reported before any GPU work."""
    import os
    from miraculix_b200 import api
    kw = {}
    if case == "magic":
        kw["magic"] = bytes([0x6C, 0x1B, 0x00])
    if case == "large":
        kw["extra"] = b"\x00"
    path, plink, Z = _triple(tmp_path, 20, 9, **kw)
    if case == "suffix":
        os.rename(path, path[:-4] + ".bin")
        path = path[:-4] + ".bin"
    if case == "nofam":
        os.remove(path[:-4] + ".fam")
    if case == "small":
        with open(path, "r+b") as fh:
            fh.truncate(3 + 20 * 3 - 1)
    with pytest.raises(api.MiraculixError):
        api.bed_shape(path)
    assert lib.mb200_last_status() == code
    with pytest.raises(api.MiraculixError):
        api.read_bed_compressed(path, 1)
    assert lib.mb200_last_status() == code
    if case not in ("suffix",):
        with pytest.raises(ValueError):
            bedfile.read_bed(path)


@pytest.mark.gpu
@pytest.mark.parametrize("snps,indiv,missing", [(1, 1, 0.0), (1003, 203, 0.02), (5000, 1025, 0.0),
                                                (300, 70001, 0.01)])
def test_read_bed_compressed_matches_oracle(gpu_lib, tmp_path, snps, indiv, missing):
    """File -> handle: device layout bit-exact against the oracle's repack of the same payload,
    frequencies = read_bed's calc_freq, and the handle multiplies like one built from memory."""
    from miraculix_b200 import api
    path, plink, Z = _triple(tmp_path, snps, indiv, seed=snps + indiv, missing=missing)
    p_o, f_o, s_o, n_o = bedfile.read_bed(path)
    h, f, s, n = api.read_bed_compressed(path, 2)
    try:
        assert (s, n) == (s_o, n_o) == (snps, indiv)
        assert np.array_equal(f, f_o)
        sm, im, cnt = api.download_layout(h)
        sm_o, im_o = cport.repack(plink, snps, indiv)
        assert np.array_equal(sm, sm_o) and np.array_equal(im, im_o)
        assert np.array_equal(cnt.astype(np.int64), cport.allele_counts(plink, snps, indiv))
        B = np.asfortranarray(np.random.default_rng(1).standard_normal((snps, 2)))
        got = api.dgemm_compressed("N", h, B, snps, indiv)
        want = cport.dgemm_compressed("N", plink, snps, indiv, f_o, B, True)
        assert np.abs(got - want).max() <= 1e-10 * max(np.abs(want).max(), 1e-300)
    finally:
        api.free_compressed(h)
