"""TEST INFRASTRUCTURE ONLY -- seeded synthetic PLINK inputs.

numpy restatement of how the reference simulates its data sets with ``plink --simulate``
(utils/genotype_simulation_plink/Makefile:22-62 and create_sim_file.jl:11: allele frequency
``rand*0.5+0.1`` per SNP), see SURVEY.md 8(d):

* ``p_j ~ U(0.1, 0.6)`` per SNP, ``z_ij ~ Binomial(2, p_j)`` i.i.d.
* PLINK codes 0 -> 00, 1 -> 10, 2 -> 11, optional missing code 01 at rate ``missing``
* SNP-major ``.bed`` payload without the 3-byte header, rows padded to whole bytes with 00
* ``f_j = sum_i z_ij / (2 indiv)`` (popcount rule of src/bindings/Julia/read_plink.jl:197-203)
"""
from __future__ import annotations

import numpy as np

_VALUE2PLINK = np.array([0b00, 0b10, 0b11], dtype=np.uint8)


def pack_2bit(codes: np.ndarray) -> np.ndarray:
    """Pack a (rows, cols) array of 2-bit codes into (rows, ceil(cols/4)) bytes, low bits first."""
    rows, cols = codes.shape
    nb = (cols + 3) // 4
    padded = np.zeros((rows, nb * 4), dtype=np.uint8)
    padded[:, :cols] = codes
    q = padded.reshape(rows, nb, 4)
    return (q[:, :, 0] | (q[:, :, 1] << 2) | (q[:, :, 2] << 4) | (q[:, :, 3] << 6)).astype(np.uint8)


def unpack_2bit(packed: np.ndarray, cols: int) -> np.ndarray:
    """Inverse of pack_2bit: (rows, nb) bytes -> (rows, cols) 2-bit codes."""
    rows, nb = packed.shape
    out = np.empty((rows, nb, 4), dtype=np.uint8)
    for t in range(4):
        out[:, :, t] = (packed >> (2 * t)) & 3
    return out.reshape(rows, nb * 4)[:, :cols]


def plink_decode(codes: np.ndarray) -> np.ndarray:
    """PLINK code -> genotype value (00->0, 01->0, 10->1, 11->2), MXinfo.h:143."""
    return np.where(codes <= 1, 0, codes.astype(np.int16) - 1).astype(np.uint8)


def simulate(snps: int, indiv: int, seed: int = 0, missing: float = 0.0):
    """Return (plink_snp_major[snps, ceil(indiv/4)] u8, Z[indiv, snps] u8 values, f[snps] f64)."""
    rng = np.random.default_rng(seed)
    p = rng.uniform(0.1, 0.6, size=snps)
    Z = rng.binomial(2, p[None, :], size=(indiv, snps)).astype(np.uint8)
    codes = _VALUE2PLINK[Z.T]  # (snps, indiv)
    if missing > 0.0:
        mask = rng.random(size=codes.shape) < missing
        codes = np.where(mask, np.uint8(0b01), codes).astype(np.uint8)
        Z = plink_decode(codes).T.copy()
    plink = pack_2bit(codes)
    f = Z.sum(axis=0, dtype=np.int64) / (2.0 * indiv)
    return np.ascontiguousarray(plink), np.ascontiguousarray(Z), f


def fortran_pattern_B(rows: int, ncol: int, scale: int = 10) -> np.ndarray:
    """B[i,j] = -(scale*i + j), 1-based -- tests/dgemm_compressed/test_5codesapi.f90:252-263
    (scale 10) and test_5codesapi_t.f90 (scale 1000).  Column-major (Fortran order)."""
    i = np.arange(1, rows + 1, dtype=np.float64)[:, None]
    j = np.arange(1, ncol + 1, dtype=np.float64)[None, :]
    return np.asfortranarray(-(scale * i + j))
