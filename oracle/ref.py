"""TEST INFRASTRUCTURE ONLY -- ctypes driver for the UNMODIFIED reference CPU "5codes" library.

``_ref/ref_miraculix.so`` is compiled by oracle/Makefile from the sources under
/root/reference/src/miraculix (that tree exists only in the build container; the prebuilt .so
travels to the GPU box).  It is driven exactly the way the reference's own benchmark drives it
(utils/benchmark/benchmark.f90:185-186): ``setOptions_compressed(0, cores, 0, 0, 1, not_center,
0, 0, 256, 0)``, ``plink2compressed(plink, plink, ...)`` (the CPU path ignores the transposed
pointer, src/miraculix/5codesChar.cc:385,393), ``dgemm_compressed``, ``free_compressed``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "_ref", "ref_miraculix.so")
REF_SRC = "/root/reference/src/miraculix"
_LIB = None


def available() -> bool:
    return os.path.exists(SO_PATH)


def build() -> bool:
    """Build _ref/ref_miraculix.so when the reference sources are present (build container)."""
    if not os.path.isdir(REF_SRC):
        return available()
    subprocess.run(["make", "-s", "-j8", "-C", _HERE, "ref"], check=True)
    return available()


def lib():
    global _LIB
    if _LIB is None:
        if not available():
            raise RuntimeError("oracle/_ref/ref_miraculix.so missing: run `make -C oracle ref` "
                               "in the container that has /root/reference")
        L = C.CDLL(SO_PATH)
        L.setOptions_compressed.argtypes = [C.c_int] * 10
        L.setOptions_compressed.restype = None
        L.plink2compressed.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                       C.POINTER(C.c_void_p)]
        L.plink2compressed.restype = None
        L.dgemm_compressed.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                       C.c_void_p, C.c_int]
        L.dgemm_compressed.restype = None
        L.free_compressed.argtypes = [C.POINTER(C.c_void_p)]
        L.free_compressed.restype = None
        _LIB = L
    return _LIB


def set_options(cores: int = 0, centered: bool = True, variant: int = 256) -> None:
    """Flag vector of utils/benchmark/benchmark.f90:186 (CPU mode, external freq)."""
    if cores <= 0:
        cores = os.cpu_count() or 1
    # silence the reference's unconditional "get started" printf (5codesAPI.c:56)
    lib()
    import sys
    sys.stdout.flush()
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    os.dup2(devnull, 1)
    try:
        lib().setOptions_compressed(0, cores, 0, 0, 1, 0 if centered else 1, 0, 0, variant, 0)
    finally:
        os.dup2(saved, 1)
        os.close(saved)
        os.close(devnull)


class RefCompressed:
    """Handle of the reference CPU path (SxI object + transposed twin)."""

    def __init__(self, plink: np.ndarray, snps: int, indiv: int, f: np.ndarray, max_n: int):
        self.plink = np.ascontiguousarray(plink, dtype=np.uint8)
        self.f = np.ascontiguousarray(f, dtype=np.float64)
        self.snps, self.indiv = snps, indiv
        self.handle = C.c_void_p(None)
        lib().plink2compressed(self.plink.ctypes.data, self.plink.ctypes.data, snps, indiv,
                               self.f.ctypes.data, max_n, C.byref(self.handle))
        if not self.handle.value:
            raise RuntimeError("reference plink2compressed returned NULL")

    def dgemm(self, trans: str, B: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        t = trans in "TtYy"
        B = np.asfortranarray(B, dtype=np.float64)
        k, n = B.shape
        m = self.snps if t else self.indiv
        assert k == (self.indiv if t else self.snps)
        if out is None:
            out = np.zeros((m, n), dtype=np.float64, order="F")
        lib().dgemm_compressed(b"T" if t else b"N", self.handle, n, B.ctypes.data, k,
                               out.ctypes.data, m)
        return out

    def free(self) -> None:
        # The reference's CPU free5 (5codesChar.cc:440-449) is flagged "will crash?!!!" by its
        # authors; leak the handle instead of risking the test process.
        self.handle = C.c_void_p(None)
