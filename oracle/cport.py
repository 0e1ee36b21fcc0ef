"""TEST INFRASTRUCTURE ONLY -- ctypes wrapper around liboracle.so (oracle.c)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build() -> str:
    """Compile oracle.c -> liboracle.so (idempotent)."""
    subprocess.run(["make", "-s", "-C", _HERE, "oracle"], check=True)
    return os.path.join(_HERE, "liboracle.so")


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        _LIB = C.CDLL(path)
    return _LIB


def _p(a, t=C.c_void_p):
    return a.ctypes.data_as(t)


def pitch128(nbytes: int) -> int:
    return ((nbytes + 127) // 128) * 128


def plink2twobit_lut() -> np.ndarray:
    t = np.zeros(256, dtype=np.uint8)
    lib().oracle_plink2twobit_lut(_p(t))
    return t


def transpose_2bit(packed: np.ndarray, rows: int, cols: int) -> np.ndarray:
    packed = np.ascontiguousarray(packed, dtype=np.uint8)
    out = np.zeros((cols, (rows + 3) // 4), dtype=np.uint8)
    lib().oracle_transpose_2bit(_p(packed), C.c_long(rows), C.c_long(cols), _p(out))
    return out


def repack(plink: np.ndarray, snps: int, indiv: int):
    plink = np.ascontiguousarray(plink, dtype=np.uint8)
    ps, pi = pitch128((indiv + 3) // 4), pitch128((snps + 3) // 4)
    sm = np.zeros((snps, ps), dtype=np.uint8)
    im = np.zeros((indiv, pi), dtype=np.uint8)
    lib().oracle_repack(_p(plink), C.c_long(snps), C.c_long(indiv), _p(sm), C.c_long(ps),
                        _p(im), C.c_long(pi))
    return sm, im


def unpack_device(packed: np.ndarray, cols: int) -> np.ndarray:
    """(rows, pitch) bytes of the device layout -> (rows, cols) 2-bit values in plain order."""
    packed = np.ascontiguousarray(packed, dtype=np.uint8)
    rows, pitch = packed.shape
    assert (cols + 15) // 16 * 4 <= pitch
    out = np.zeros((rows, cols), dtype=np.uint8)
    lib().oracle_unpack_device(_p(packed), C.c_long(rows), C.c_long(pitch), C.c_long(cols), _p(out))
    return out


def allele_counts(plink: np.ndarray, snps: int, indiv: int) -> np.ndarray:
    plink = np.ascontiguousarray(plink, dtype=np.uint8)
    out = np.zeros(snps, dtype=np.int64)
    lib().oracle_allele_counts(_p(plink), C.c_long(snps), C.c_long(indiv), _p(out))
    return out


def dgemm_compressed(trans: str, plink: np.ndarray, snps: int, indiv: int, f: np.ndarray,
                     B: np.ndarray, centered: bool = True) -> np.ndarray:
    t = 1 if trans in "TtYy" else 0
    plink = np.ascontiguousarray(plink, dtype=np.uint8)
    f = np.ascontiguousarray(f, dtype=np.float64)
    B = np.asfortranarray(B, dtype=np.float64)
    k, n = B.shape
    m = snps if t else indiv
    assert k == (indiv if t else snps)
    Cm = np.zeros((m, n), dtype=np.float64, order="F")
    lib().oracle_dgemm_compressed(C.c_int(t), _p(plink), C.c_long(snps), C.c_long(indiv), _p(f),
                                  C.c_int(int(centered)), C.c_int(n), _p(B), C.c_long(k), _p(Cm),
                                  C.c_long(m))
    return Cm


def crossprod(x: np.ndarray, k: int, rows: int, is_plink_format: bool = False) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.uint8)
    ans = np.zeros((rows, rows), dtype=np.float64, order="F")
    lib().oracle_crossprod(_p(x), C.c_long(k), C.c_long(rows), _p(ans), C.c_int(int(is_plink_format)))
    return ans


def grm_epilogue(M: np.ndarray, f: np.ndarray, do_scale: bool = True) -> np.ndarray:
    M = np.array(M, dtype=np.float64, order="F", copy=True)
    f = np.ascontiguousarray(f, dtype=np.float64)
    lib().oracle_grm_epilogue(_p(M), C.c_long(M.shape[0]), _p(f), C.c_long(f.shape[0]),
                              C.c_int(int(do_scale)))
    return M


def ld_epilogue(M: np.ndarray, n_indiv: int, f: np.ndarray) -> np.ndarray:
    M = np.array(M, dtype=np.float64, order="F", copy=True)
    f = np.ascontiguousarray(f, dtype=np.float64)
    lib().oracle_ld_epilogue(_p(M), C.c_long(M.shape[0]), C.c_long(n_indiv), _p(f))
    return M
