"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's read_bed
(src/bindings/Julia/read_plink.jl:161-228) and a writer of PLINK .bed/.bim/.fam triples for the
tests (the layout read_bed documents: magic 6c 1b 01, then one block of ceil(indiv/4) bytes per SNP)."""
from __future__ import annotations

import os

import numpy as np

MAGIC = bytes([0x6C, 0x1B, 0x01])


def write_bed(stem: str, plink: np.ndarray, snps: int, indiv: int, magic: bytes = MAGIC,
              extra: bytes = b"") -> str:
    plink = np.ascontiguousarray(plink, dtype=np.uint8)
    assert plink.shape == (snps, (indiv + 3) // 4)
    with open(stem + ".bed", "wb") as fh:
        fh.write(magic)
        fh.write(plink.tobytes())
        fh.write(extra)
    with open(stem + ".bim", "w") as fh:
        for j in range(snps):
            fh.write(f"1\tsnp{j}\t0\t{j + 1}\tA\tC\n")
    with open(stem + ".fam", "w") as fh:
        for i in range(indiv):
            fh.write(f"fam{i} ind{i} 0 0 0 -9\n")
    return stem + ".bed"


def read_bed(file: str):
    """(plink[snps, ceil(indiv/4)], freq[snps], n_snps, n_indiv); errors as read_bed raises them."""
    if not file.endswith(".bed"):                                              # :163-165
        raise ValueError("File not in .bed format")
    stem = file[:-4]
    if not (os.path.isfile(stem + ".fam") and os.path.isfile(stem + ".bim")):  # :166-168
        raise ValueError("Missing supplementary file .fam or .bim")
    raw = open(file, "rb").read()
    if raw[:3] != MAGIC:                                                       # :175-177
        raise ValueError("Not a correct .bed file")
    n_indiv = open(stem + ".fam", "rb").read().count(b"\n")                    # :181-182
    n_snps = open(stem + ".bim", "rb").read().count(b"\n")
    nb = (n_indiv + 3) // 4
    if len(raw) < 3 + n_snps * nb:
        raise ValueError("Too small .bed file")                               # unsafe_read throws
    if len(raw) > 3 + n_snps * nb:
        raise ValueError("Too large .bed file")                               # :193
    plink = np.frombuffer(raw, dtype=np.uint8, offset=3).reshape(n_snps, nb).copy()
    ones = np.unpackbits(plink, axis=1).sum(axis=1, dtype=np.int64)            # count_ones, :199-203
    return plink, ones / (2.0 * n_indiv), n_snps, n_indiv
