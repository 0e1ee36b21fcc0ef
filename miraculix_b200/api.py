"""Host-side mirror of the reference's Julia bindings over the C-ABI of libmiraculix_b200.so.

Function names and argument meaning follow
  src/bindings/Julia/dgemm_compressed.jl:42-155  (set_options, init_compressed, dgemm_compressed, free_compressed)
  src/bindings/Julia/crossproduct.jl:44-152      (snp_crossprod, grm, ld)
of the reference so that tests read like tests/dgemm_compressed/test.jl and
tests/crossproduct/test_grm.jl.  Everything here is ctypes plumbing: no arithmetic happens in
Python, and there is no CPU fallback -- a missing library or GPU raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmiraculix_b200.so")
_LIB = None


class Layout(C.Structure):
    _fields_ = [("snps", C.c_long), ("indiv", C.c_long), ("pitch_snp_major", C.c_long),
                ("pitch_indiv_major", C.c_long), ("d_snp_major", C.c_void_p),
                ("d_indiv_major", C.c_void_p), ("d_f", C.c_void_p), ("d_allele_counts", C.c_void_p),
                ("device", C.c_int)]


# every symbol include/miraculix_b200.h declares: (restype, argtypes)
SIGNATURES = {
    "setOptions_compressed": (None, [C.c_int] * 10),
    "plink2compressed": (None, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                C.POINTER(C.c_void_p)]),
    "dgemm_compressed": (None, [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p,
                                C.c_int]),
    "free_compressed": (None, [C.POINTER(C.c_void_p)]),
    "get_compressed_freq": (None, [C.c_void_p, C.c_void_p]),
    "snp_multiply_gpu": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_bool]),
    "mb200_last_status": (C.c_int, []),
    "mb200_last_message": (C.c_char_p, []),
    "mb200_kernel_launches": (C.c_ulonglong, []),
    "mb200_set_dgemm_mode": (C.c_int, [C.c_int, C.c_int]),
    "mb200_set_crossprod_mode": (C.c_int, [C.c_int]),
    "mb200_set_dgemm_small_limit": (C.c_int, [C.c_double]),
    "mb200_set_dgemm_wide_cluster": (C.c_int, [C.c_int]),
    "mb200_set_pipeline": (C.c_int, [C.c_int, C.c_double]),
    "mb200_dgemm_compressed_device": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_long,
                                                C.c_void_p, C.c_long]),
    "mb200_dgemm_compressed_noalloc": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_long,
                                                 C.c_void_p, C.c_long]),
    "mb200_noalloc_max_cols": (C.c_int, []),
    "mb200_dgemm_compressed_mixed": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_long,
                                               C.c_int, C.c_void_p, C.c_long, C.c_int]),
    "mb200_stream": (C.c_void_p, [C.c_void_p]),
    "mb200_debug_parallel_copy": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t,
                                            C.c_size_t]),
    "mb200_sync": (C.c_int, [C.c_void_p]),
    "mb200_dgemm_last_precision": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "mb200_debug_read_trace": (C.c_int, [C.POINTER(C.c_longlong), C.c_int]),
    "mb200_dgemm_last_kernel_ms": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "mb200_dgemm_kernel_times": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float),
                                           C.POINTER(C.c_int)]),
    "mb200_plink2compressed_device": (C.c_int, [C.c_void_p, C.c_long, C.c_long, C.c_void_p, C.c_int,
                                                C.POINTER(C.c_void_p)]),
    "mb200_get_layout": (C.c_int, [C.c_void_p, C.POINTER(Layout)]),
    "mb200_download_layout": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "mb200_simulate_plink_device": (C.c_int, [C.c_void_p, C.c_long, C.c_long, C.c_ulonglong]),
    "mb200_simulate_plink_rows_device": (C.c_int, [C.c_void_p, C.c_long, C.c_long, C.c_ulonglong, C.c_long]),
    "mb200_crossprod_device": (C.c_int, [C.c_void_p, C.c_long, C.c_long, C.c_int, C.c_int,
                                         C.c_void_p, C.c_int, C.c_void_p]),
    "mb200_grm": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_bool, C.c_void_p, C.c_bool]),
    "mb200_ld": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_bool, C.c_void_p]),
    "mb200_crossprod_partition_rows": (C.c_long, [C.c_long, C.c_int, C.c_int]),
    "mb200_crossprod_partition_device": (C.c_int, [C.c_void_p, C.c_long, C.c_long, C.c_int, C.c_int,
                                                   C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "mb200_grm_partition": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_bool, C.c_int, C.c_void_p,
                                      C.c_bool, C.c_int, C.c_int, C.c_void_p]),
    "mb200_crossprod_last_timing": (C.c_int, [C.POINTER(C.c_double * 3)]),
    "sparse_times_plink": (None, [C.c_char_p, C.c_char_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int,
                                  C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]),
    "dgemm_plink": (None, [C.c_char_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int,
                           C.c_void_p, C.c_int, C.c_void_p, C.c_int]),
    "potrs_solve_gpu": (None, [C.c_void_p, C.c_uint, C.c_void_p, C.c_uint, C.c_void_p, C.c_void_p, C.c_int,
                               C.POINTER(C.c_int)]),
    "mb200_potrs_solve_device": (C.c_int, [C.c_void_p, C.c_long, C.c_void_p, C.c_long, C.c_void_p]),
    "mb200_bed_shape": (C.c_int, [C.c_char_p, C.POINTER(C.c_long), C.POINTER(C.c_long)]),
    "mb200_read_bed": (C.c_int, [C.c_char_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_long),
                                 C.POINTER(C.c_long)]),
    "mb200_cg_solve_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double,
                                        C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_int),
                                        C.POINTER(C.c_double)]),
    "mb200_set_devices": (C.c_int, [C.c_int, C.POINTER(C.c_int)]),
    "mb200_shard_count": (C.c_int, [C.c_void_p]),
    "mb200_get_shard": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_long),
                                  C.POINTER(C.c_long)]),
    "mb200_simulate_compressed": (C.c_int, [C.c_long, C.c_long, C.c_ulonglong, C.c_int,
                                            C.POINTER(C.c_void_p)]),
    "mb200_cg_solve": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int,
                                 C.POINTER(C.c_int), C.POINTER(C.c_double)]),
}
# mb200_allreduce_fn: int (*)(double *d_buf, long count, void *stream, void *user)
ALLREDUCE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p)


def load_library(path: str | None = None):
    """dlopen the C-ABI library (cf. load_shared_library, src/bindings/Julia/miraculix.jl)."""
    global _LIB
    if _LIB is not None and path is None:
        return _LIB
    path = path or os.environ.get("MB200_LIB_PATH") or LIB_PATH  # env override: A/B kernel builds
    if not os.path.exists(path):
        raise RuntimeError(f"{path} not built: run `python -m miraculix_b200.build` "
                           "(there is no CPU fallback)")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype, fn.argtypes = res, args
    _LIB = lib
    return lib


class MiraculixError(RuntimeError):
    pass


def _check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load_library().mb200_last_message().decode(errors="replace")
        raise MiraculixError(f"{what} failed ({rc}): {msg}")


def last_status() -> int:
    return load_library().mb200_last_status()


def kernel_launches() -> int:
    return int(load_library().mb200_kernel_launches())


def set_options(use_gpu: bool = True, cores: int = 0, not_center: bool = False, variant: int = 0,
                verbose: int = 0) -> None:
    """dgemm_compressed.jl:42-58 -- same flag vector: floatLoop=0, meanSubstract=0,
    ignore_missings=1, do_normalize=0, use_miraculix_freq=0."""
    load_library().setOptions_compressed(int(use_gpu), cores, 0, 0, 1, int(not_center), 0, 0,
                                         variant, verbose)


def init_compressed(plink: np.ndarray, snps: int, indiv: int, freq: np.ndarray, max_ncol: int,
                    plink_transposed: np.ndarray | None = None) -> C.c_void_p:
    """dgemm_compressed.jl:81-92.  ``plink``: (snps, ceil(indiv/4)) uint8 rows = SNPs (the Julia
    matrix is (ceil(indiv/4), snps) column-major: same memory).  ``plink_transposed`` is accepted
    and ignored -- the GPU repack produces both orientations."""
    plink = np.ascontiguousarray(plink, dtype=np.uint8)
    freq = np.ascontiguousarray(freq, dtype=np.float64)
    if plink.shape != (snps, (indiv + 3) // 4):
        raise ValueError(f"plink has shape {plink.shape}, expected {(snps, (indiv + 3) // 4)}")
    if freq.shape != (snps,):
        raise ValueError("freq must have length snps")
    handle = C.c_void_p(None)
    tp = plink_transposed.ctypes.data if plink_transposed is not None else plink.ctypes.data
    load_library().plink2compressed(plink.ctypes.data, tp, snps, indiv, freq.ctypes.data, max_ncol,
                                    C.byref(handle))
    if not handle.value:  # dgemm_compressed.jl:90
        raise MiraculixError("plink2compressed returned NULL: " +
                             load_library().mb200_last_message().decode(errors="replace"))
    return handle


def dgemm_compressed(trans: str, handle: C.c_void_p, B: np.ndarray, snps: int, indiv: int,
                     out: np.ndarray | None = None) -> np.ndarray:
    """dgemm_compressed.jl:115-134: 'N' -> (indiv x n) = (Z - 2 1 f^T) B, 'T' -> (snps x n)."""
    t = trans in "TtYy"
    B = np.asfortranarray(B, dtype=np.float64)
    if B.ndim == 1:
        B = B.reshape(-1, 1, order="F")
    k, n = B.shape
    m = snps if t else indiv
    if k != (indiv if t else snps):
        raise ValueError(f"B has {k} rows, expected {indiv if t else snps}")
    if out is None:
        out = np.zeros((m, n), dtype=np.float64, order="F")
    lib = load_library()
    lib.dgemm_compressed(trans.encode()[:1], handle, n, B.ctypes.data, k, out.ctypes.data, m)
    _check(lib.mb200_last_status(), "dgemm_compressed")
    return out


def set_devices(devices=None) -> None:
    """Devices the next handles / crossproducts use (MIRACULIX_NGPUS / CUDA_DEVICE=a,b,c in code);
    None or [] = back to the environment."""
    devices = list(devices or [])
    arr = (C.c_int * max(1, len(devices)))(*devices)
    _check(load_library().mb200_set_devices(len(devices), arr), "mb200_set_devices")


def shards(handle: C.c_void_p):
    """[(shard_handle, snp_begin, snp_count)] of a (possibly multi-device) handle."""
    lib = load_library()
    out = []
    for g in range(lib.mb200_shard_count(handle)):
        sh, b, c = C.c_void_p(None), C.c_long(0), C.c_long(0)
        _check(lib.mb200_get_shard(handle, g, C.byref(sh), C.byref(b), C.byref(c)), "mb200_get_shard")
        out.append((sh, b.value, c.value))
    return out


def simulate_compressed(snps: int, indiv: int, seed: int, max_ncol: int) -> C.c_void_p:
    """Handle from the on-device simulator (one device, or sharded per set_devices / environment)."""
    h = C.c_void_p(None)
    _check(load_library().mb200_simulate_compressed(snps, indiv, seed, max_ncol, C.byref(h)),
           "mb200_simulate_compressed")
    return h


def free_compressed(handle: C.c_void_p) -> None:
    """dgemm_compressed.jl:150-155."""
    load_library().free_compressed(C.byref(handle))


def get_compressed_freq(handle: C.c_void_p, snps: int) -> np.ndarray:
    f = np.zeros(snps, dtype=np.float64)
    load_library().get_compressed_freq(handle, f.ctypes.data)
    return f


def get_layout(handle: C.c_void_p) -> Layout:
    lay = Layout()
    _check(load_library().mb200_get_layout(handle, C.byref(lay)), "mb200_get_layout")
    return lay


def download_layout(handle: C.c_void_p):
    """(snp_major[snps, pitch_s], indiv_major[indiv, pitch_i], allele_counts[snps]) from HBM."""
    lay = get_layout(handle)
    sm = np.zeros((lay.snps, lay.pitch_snp_major), dtype=np.uint8)
    im = np.zeros((lay.indiv, lay.pitch_indiv_major), dtype=np.uint8)
    cnt = np.zeros(lay.snps, dtype=np.uint64)
    _check(load_library().mb200_download_layout(handle, sm.ctypes.data, im.ctypes.data,
                                                cnt.ctypes.data), "mb200_download_layout")
    return sm, im, cnt


def _check_packed(x: np.ndarray, k: int, rows: int) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.uint8)
    if x.shape != (rows, (k + 3) // 4):  # check_dimensions, crossproduct.jl:51
        raise ValueError(f"matrix has shape {x.shape}, expected {(rows, (k + 3) // 4)}")
    return x


def snp_crossprod(x: np.ndarray, snps: int, indiv: int, is_snpmajor: bool,
                  is_plink_format: bool = False) -> np.ndarray:
    """crossproduct.jl:44-64.  is_snpmajor=False: x = (indiv, ceil(snps/4)) -> Z Z^T (indiv^2);
    is_snpmajor=True: x = (snps, ceil(indiv/4)) -> Z^T Z (snps^2)."""
    nrow, ncol = (indiv, snps) if is_snpmajor else (snps, indiv)
    x = _check_packed(x, nrow, ncol)
    M = np.zeros((ncol, ncol), dtype=np.float64, order="F")
    rc = load_library().snp_multiply_gpu(x.ctypes.data, nrow, ncol, M.ctypes.data, is_plink_format)
    _check(rc, "snp_multiply_gpu")
    return M


def grm(x: np.ndarray, snps: int, indiv: int, is_plink_format: bool = False, do_scale: bool = True,
        allele_freq: np.ndarray | None = None) -> np.ndarray:
    """crossproduct.jl:83-110, with the centring/scaling epilogue fused on the GPU."""
    if do_scale and (allele_freq is None or len(allele_freq) != snps):
        raise ValueError(f"Allele frequencies need to be equal to length of SNPs {snps}.")
    x = _check_packed(x, snps, indiv)
    f = np.ascontiguousarray(allele_freq, dtype=np.float64) if allele_freq is not None else None
    M = np.zeros((indiv, indiv), dtype=np.float64, order="F")
    rc = load_library().mb200_grm(x.ctypes.data, snps, indiv, M.ctypes.data, is_plink_format,
                                  f.ctypes.data if f is not None else None, do_scale)
    _check(rc, "mb200_grm")
    return M


def ld(x: np.ndarray, snps: int, indiv: int, is_plink_format: bool = False,
       allele_freq: np.ndarray | None = None) -> np.ndarray:
    """crossproduct.jl:128-152: x SNP-major (snps, ceil(indiv/4))."""
    if allele_freq is None or len(allele_freq) != snps:
        raise ValueError(f"Allele frequencies need to be equal to length of SNPs {snps}.")
    x = _check_packed(x, indiv, snps)
    f = np.ascontiguousarray(allele_freq, dtype=np.float64)
    M = np.zeros((snps, snps), dtype=np.float64, order="F")
    rc = load_library().mb200_ld(x.ctypes.data, indiv, snps, M.ctypes.data, is_plink_format,
                                 f.ctypes.data)
    _check(rc, "mb200_ld")
    return M


def grm_partition(x: np.ndarray, snps: int, indiv: int, part: int, nparts: int, mode: int = 1,
                  is_plink_format: bool = False, do_scale: bool = True,
                  allele_freq: np.ndarray | None = None):
    """One partition's panel of the (centred) GRM: (global_row_index[local_rows], panel[local_rows,
    indiv]); entries left of the diagonal tile are 0 (see include/miraculix_b200.h)."""
    x = _check_packed(x, snps, indiv)
    lib = load_library()
    lrows = lib.mb200_crossprod_partition_rows(indiv, part, nparts)
    panel = np.zeros((lrows, indiv), dtype=np.float64)
    f = np.ascontiguousarray(allele_freq, dtype=np.float64) if allele_freq is not None else None
    if lrows:
        rc = lib.mb200_grm_partition(x.ctypes.data, snps, indiv, is_plink_format, mode,
                                     f.ctypes.data if f is not None else None, do_scale, part, nparts,
                                     panel.ctypes.data)
        _check(rc, "mb200_grm_partition")
    from .sharded import partition_tile_rows
    return partition_tile_rows(indiv, part, nparts), panel


def dense_solve(M: np.ndarray, B: np.ndarray, calc_logdet: bool = False, oversubscribe: bool = False):
    """src/bindings/Julia/solve.jl:139-180: X = M \\ B for symmetric positive definite M through
    potrs_solve_gpu; with calc_logdet returns (X, logdet)."""
    M = np.asfortranarray(M, dtype=np.float64)
    B2 = np.asfortranarray(B, dtype=np.float64).reshape(M.shape[0], -1, order="F")
    if M.shape[0] != M.shape[1]:
        raise ValueError("M must be square")
    X = np.zeros_like(B2, order="F")
    logdet, status = C.c_double(0.0), C.c_int(1)
    load_library().potrs_solve_gpu(M.ctypes.data, M.shape[0], B2.ctypes.data, B2.shape[1], X.ctypes.data,
                                   C.cast(C.byref(logdet), C.c_void_p) if calc_logdet else None,
                                   int(oversubscribe), C.byref(status))
    if status.value != 0:
        raise MiraculixError("potrs_solve_gpu failed: " +
                             load_library().mb200_last_message().decode(errors="replace"))
    X = X.reshape(np.shape(B), order="F")
    return (X, logdet.value) if calc_logdet else X


def sparse_times_plink(transcompressed: str, plink: np.ndarray, plink_transposed, snps: int,
                       indiv: int, row_idx: np.ndarray, col_idx: np.ndarray, values: np.ndarray,
                       ldc: int | None = None) -> np.ndarray:
    """c_sparse_times_plink of mod5codesapi.f90:84-100: CSR (0-based) times PLINK genotypes.
    'N': plink (snps, ceil(indiv/4)) -> (nIdx, indiv); 'T': plink_transposed (indiv, ceil(snps/4))
    -> (nIdx, snps).  Returns the (ldc, columns) column-major array."""
    t = transcompressed in "TtYy"
    row_idx = np.ascontiguousarray(row_idx, dtype=np.int32)
    col_idx = np.ascontiguousarray(col_idx, dtype=np.int32)
    values = np.ascontiguousarray(values, dtype=np.float64)
    n_idx = len(row_idx) - 1
    ldc = n_idx if ldc is None else ldc
    code = np.ascontiguousarray(plink_transposed if t else plink, dtype=np.uint8)
    width = snps if t else indiv
    out = np.zeros((ldc, width), dtype=np.float64, order="F")
    lib = load_library()
    lib.sparse_times_plink(b"N", transcompressed.encode()[:1], code.ctypes.data, code.ctypes.data, snps,
                           indiv, n_idx, row_idx.ctypes.data, col_idx.ctypes.data, values.ctypes.data,
                           out.ctypes.data, ldc)
    _check(lib.mb200_last_status(), "sparse_times_plink")
    return out


def dgemm_plink(trans: str, plink: np.ndarray, snps: int, indiv: int, B: np.ndarray,
                freq: np.ndarray | None = None) -> np.ndarray:
    """One-shot product straight from the bed payload (5codesAPI.c:112-127)."""
    t = trans in "TtYy"
    plink = np.ascontiguousarray(plink, dtype=np.uint8)
    B = np.asfortranarray(B, dtype=np.float64)
    k, n = B.shape
    m = snps if t else indiv
    out = np.zeros((m, n), dtype=np.float64, order="F")
    f = np.ascontiguousarray(freq, dtype=np.float64) if freq is not None else None
    lib = load_library()
    lib.dgemm_plink(trans.encode()[:1], plink.ctypes.data, plink.ctypes.data, snps, indiv,
                    f.ctypes.data if f is not None else None, n, B.ctypes.data, k, out.ctypes.data, m)
    _check(lib.mb200_last_status(), "dgemm_plink")
    return out


def bed_shape(file: str):
    """(n_snps, n_indiv) of a PLINK .bed/.bim/.fam triple after read_bed's checks
    (read_plink.jl:161-195); raises MiraculixError like read_bed's error() calls.  No GPU needed."""
    s, n = C.c_long(0), C.c_long(0)
    _check(load_library().mb200_bed_shape(file.encode(), C.byref(s), C.byref(n)), "mb200_bed_shape")
    return s.value, n.value


def read_bed_compressed(file: str, max_ncol: int):
    """read_bed(file, calc_freq=true) + transpose_genotype_matrix + init_compressed in one call:
    the file is streamed to the GPU and repacked there.  Returns (handle, freq, n_snps, n_indiv)."""
    h, s, n = C.c_void_p(None), C.c_long(0), C.c_long(0)
    _check(load_library().mb200_read_bed(file.encode(), max_ncol, C.byref(h), C.byref(s), C.byref(n)),
           "mb200_read_bed")
    return h, get_compressed_freq(h, s.value), s.value, n.value


def cg_solve(handle: C.c_void_p, b: np.ndarray, x0: np.ndarray, lam: float = 0.0,
             tol: float = 1e-2, max_iter: int = 1000):
    """The conjugate-gradient loop of examples/iterative_solver/grm_solve_cg.jl:109-133 run on the
    device: solves (Zc Zc^T + lam I) x = b from the start vector x0 (host vectors of length
    indiv).  Returns (x, iterations, ||r||)."""
    b = np.ascontiguousarray(b, dtype=np.float64).ravel()
    x = np.array(x0, dtype=np.float64, copy=True).ravel()
    if b.shape != x.shape:
        raise ValueError("b and x0 must have the same length (indiv)")
    it, res = C.c_int(0), C.c_double(0.0)
    rc = load_library().mb200_cg_solve(handle, b.ctypes.data, x.ctypes.data, lam, tol, max_iter,
                                       C.byref(it), C.byref(res))
    _check(rc, "mb200_cg_solve")
    return x, it.value, res.value


def cg_solve_device(handle: C.c_void_p, d_b: int, d_x: int, lam: float, tol: float, max_iter: int,
                    allreduce=None):
    """Device-pointer variant; `allreduce` is an ALLREDUCE_FN (or None) summing the rank's partial
    matrix-vector product over the SNP shards (see include/miraculix_b200.h)."""
    it, res = C.c_int(0), C.c_double(0.0)
    cb = C.cast(allreduce, C.c_void_p) if allreduce is not None else None
    rc = load_library().mb200_cg_solve_device(handle, C.c_void_p(d_b), C.c_void_p(d_x), lam, tol,
                                              max_iter, cb, None, C.byref(it), C.byref(res))
    _check(rc, "mb200_cg_solve_device")
    return it.value, res.value


def crossprod_last_timing():
    out = (C.c_double * 3)()
    load_library().mb200_crossprod_last_timing(C.byref(out))
    return {"gemm_ms": out[0], "epilogue_ms": out[1], "executed_ops": out[2]}
