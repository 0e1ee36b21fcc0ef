"""SNP-block sharding of dgemm_compressed over one process per GPU.

The reference is single-GPU (src/cuda/dgemm_compressed_cuda.cu:43-170 uploads the whole matrix to
one device).  Here rank g owns the SNP range S_g (boundaries multiples of 512 SNPs = one 128-byte
line of the individual-major layout) in BOTH device orientations, so every direction streams each
rank's shard exactly once:

  'T'  C[S_g, :] = (Z[:, S_g] - 2 1 f_g^T)^T B          no communication, outputs are disjoint
  'N'  C = sum_g (Z[:, S_g] - 2 1 f_g^T) B[S_g, :]      ONE all-reduce (sum, fp64) of indiv x n

The centring term is linear in the SNP range, so each rank folds its own share into its partial
before the reduce.  `torch.distributed` is the plumbing (NCCL on GPUs; gloo in the CPU tests,
which inject the oracle as the local product because the library has no CPU path).
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, List, Optional, Tuple

BLOCK = 512  # SNPs per 128-byte line of the individual-major rows


def partition_snps(snps: int, world: int) -> List[Tuple[int, int]]:
    """(begin, count) per rank: contiguous, multiples of 512 SNPs except the tail, balanced to
    within one block.  Ranks beyond the number of blocks get an empty range."""
    if snps < 0 or world < 1:
        raise ValueError("snps >= 0 and world >= 1 required")
    nblocks = (snps + BLOCK - 1) // BLOCK
    base, extra = divmod(nblocks, world)
    out, begin = [], 0
    for r in range(world):
        blocks = base + (1 if r < extra else 0)
        end = min(snps, begin + blocks * BLOCK)
        out.append((begin, end - begin))
        begin = end
    assert begin == snps
    return out


TILE = 128  # rows per output tile of the crossproduct partition (csrc/crossprod.cu)


def partition_tile_rows(rows: int, part: int, nparts: int):
    """Global row index of every local row of partition `part` of `nparts` of the 2-D output-tile
    partition of the GRM (include/miraculix_b200.h, mb200_crossprod_partition_device): the
    partition owns the 128-row tiles t with t % nparts == part; indices >= rows are the padding of
    the last tile.  Same length as mb200_crossprod_partition_rows(rows, part, nparts)."""
    import numpy as np
    if rows < 0 or nparts < 1 or not 0 <= part < nparts:
        raise ValueError("rows >= 0 and 0 <= part < nparts required")
    tiles = (rows + TILE - 1) // TILE
    owned = (tiles - part + nparts - 1) // nparts if part < tiles else 0
    lr = np.arange(owned * TILE, dtype=np.int64)
    return ((lr // TILE) * nparts + part) * TILE + lr % TILE


def assemble_grm_panels(panels, rows: int):
    """Full symmetric rows x rows matrix from the panels of ALL partitions (panels[part] =
    [local_rows][rows], row-major, as mb200_crossprod_partition_device writes them).  A panel row
    holds global row i from column 256 floor(i/256) on; the rest comes from symmetry -- the
    reference fills the lower triangle the same way (crossproduct.jl:100-106)."""
    import numpy as np
    nparts = len(panels)
    out = np.zeros((rows, rows), dtype=np.float64)
    seen = np.zeros(rows, dtype=bool)
    for part, panel in enumerate(panels):
        idx = partition_tile_rows(rows, part, nparts)
        panel = np.asarray(panel)
        if panel.shape != (idx.size, rows):
            raise ValueError(f"panel {part}: shape {panel.shape}, expected {(idx.size, rows)}")
        for lr in np.flatnonzero(idx < rows):
            i = int(idx[lr])
            first = (i // 256) * 256
            out[i, first:] = panel[lr, first:]
            seen[i] = True
    if not seen.all():
        raise ValueError("panels do not cover every row")
    iu = np.triu_indices(rows, 1)
    out.T[iu] = out[iu]
    return out


class ShardedCompressed:
    """Rank-local view of a SNP-block-sharded compressed matrix.

    handle       C-ABI handle of THIS rank's shard (snps_local x indiv), or None when a
                 `local_dgemm` callable is injected (tests).
    local_dgemm  callable(trans:int, B_local, C_out) computing the rank-local centred product on
                 tensors; defaults to mb200_dgemm_compressed_device on the handle's stream.
    Panels: a column-major rows x n matrix is an (n, rows) row-major contiguous torch tensor.
    """

    def __init__(self, handle, snps_total: int, indiv: int, rank: int, world: int,
                 group=None, local_dgemm: Optional[Callable] = None):
        self.handle, self.snps_total, self.indiv = handle, snps_total, indiv
        self.rank, self.world, self.group = rank, world, group
        self.parts = partition_snps(snps_total, world)
        self.s_begin, self.s_count = self.parts[rank]
        self._local = local_dgemm or self._device_dgemm
        self._stream = None

    # -- device plumbing --------------------------------------------------------------------
    def stream(self):
        """The library's stream as a torch stream (so NCCL and events order against it)."""
        import torch
        from . import api
        if self._stream is None:
            ptr = api.load_library().mb200_stream(self.handle)
            self._stream = torch.cuda.ExternalStream(ptr)
        return self._stream

    def _fork(self):
        """Order the library's stream after the caller's current torch stream (inputs produced
        there are complete before our kernels read them)."""
        import torch
        if self.handle is not None and torch.cuda.is_available():
            self.stream().wait_stream(torch.cuda.current_stream())

    def _join(self):
        """... and the caller's stream after ours (later torch ops see finished results)."""
        import torch
        if self.handle is not None and torch.cuda.is_available():
            torch.cuda.current_stream().wait_stream(self.stream())

    def _device_dgemm(self, trans: int, B, Cout) -> None:
        from . import api
        lib = api.load_library()
        # a column-major rows x n panel is held as an (n, rows) row-major contiguous tensor
        n = B.shape[0]
        rc = lib.mb200_dgemm_compressed_device(trans, self.handle, n, C.c_void_p(B.data_ptr()),
                                               B.shape[1], C.c_void_p(Cout.data_ptr()),
                                               Cout.shape[1])
        api._check(rc, "mb200_dgemm_compressed_device")

    # -- the two directions -----------------------------------------------------------------
    def dgemm_n(self, B_local, C_out):
        """B_local: (n, |S_g|) this rank's rows of B; C_out: (n, indiv) -- the full product."""
        import torch.distributed as dist
        cuda = bool(getattr(C_out, "is_cuda", False))
        if cuda:
            self._fork()
        if self.s_count > 0:
            self._local(0, B_local, C_out)
        elif cuda and self.handle is not None:
            import torch
            with torch.cuda.stream(self.stream()):
                C_out.zero_()
        else:
            C_out.zero_()
        if self.world > 1:
            if cuda and self.handle is not None:
                import torch
                with torch.cuda.stream(self.stream()):
                    dist.all_reduce(C_out, op=dist.ReduceOp.SUM, group=self.group)
            else:
                dist.all_reduce(C_out, op=dist.ReduceOp.SUM, group=self.group)
        if cuda:
            self._join()
        return C_out

    def dgemm_t(self, B, C_local):
        """B: (n, indiv) replicated; C_local: (n, |S_g|) rows S_g of the result.  No exchange."""
        cuda = bool(getattr(C_local, "is_cuda", False))
        if cuda:
            self._fork()
        if self.s_count > 0:
            self._local(1, B, C_local)
        if cuda:
            self._join()
        return C_local

    # -- the same two directions with HOST panels (pinned CPU tensors) -------------------------
    def dgemm_n_host(self, hB_local, C_dev, hC):
        """'N' end to end: hB_local (n, |S_g|) host rows of B of this rank -> hC (n, indiv) host.
        The upload is pipelined against the kernels inside the library
        (mb200_dgemm_compressed_mixed: B on the host, partial C on the device), the partial is
        all-reduced on the device, then downloaded asynchronously on the library's stream (the
        next blocking call or stream().synchronize() completes it)."""
        import torch
        import torch.distributed as dist
        from . import api
        lib = api.load_library()
        n = hB_local.shape[0]
        self._fork()
        if self.s_count > 0:
            rc = lib.mb200_dgemm_compressed_mixed(0, self.handle, n, C.c_void_p(hB_local.data_ptr()),
                                                  hB_local.shape[1], 0, C.c_void_p(C_dev.data_ptr()),
                                                  C_dev.shape[1], 1)
            api._check(rc, "mb200_dgemm_compressed_mixed")
        else:
            C_dev.zero_()
        with torch.cuda.stream(self.stream()):
            if self.world > 1:
                dist.all_reduce(C_dev, op=dist.ReduceOp.SUM, group=self.group)
            hC.copy_(C_dev, non_blocking=True)
        self._join()
        return hC

    def dgemm_t_host(self, hB, hC_local):
        """'T' end to end: hB (n, indiv) host, replicated -> hC_local (n, |S_g|) host rows S_g.
        The reference-facing dgemm_compressed() itself on this rank's shard: no exchange, the
        download of C is pipelined against the kernels.  Blocking."""
        from . import api
        lib = api.load_library()
        if self.s_count > 0:
            lib.dgemm_compressed(b"T", self.handle, hB.shape[0], hB.data_ptr(), hB.shape[1],
                                 hC_local.data_ptr(), hC_local.shape[1])
            api._check(lib.mb200_last_status(), "dgemm_compressed")
        return hC_local

    # -- conjugate gradients over the shards ------------------------------------------------
    def cg_solve(self, b, x, lam: float = 0.0, tol: float = 1e-2, max_iter: int = 1000):
        """(Zc Zc^T + lam I) x = b with the matrix SNP-sharded over the ranks (BASELINE
        configs[3]; loop of examples/iterative_solver/grm_solve_cg.jl:109-133 inside the C-ABI
        library, mb200_cg_solve_device).  b, x: replicated CUDA tensors of length indiv; x holds
        the start vector and receives the iterate.  The only exchange is ONE all-reduce of indiv
        doubles per matrix-vector product, issued from the library's hook on its own stream.
        Returns (iterations, ||r||)."""
        import torch
        import torch.distributed as dist
        from . import api

        hook = None
        if self.world > 1:
            stream, group, dev = self.stream(), self.group, x.device

            class _Dev:  # a device buffer of the library seen through __cuda_array_interface__
                def __init__(self, ptr, n):
                    self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8",
                                                     "data": (ptr, False), "version": 2}

            def _allreduce(ptr, count, _stream, _user):
                try:
                    t = torch.as_tensor(_Dev(ptr, count), device=dev)
                    with torch.cuda.stream(stream):
                        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
                    return 0
                except Exception as exc:  # never unwind through the C frames
                    print(f"sharded cg: all-reduce failed: {exc}")
                    return 1

            hook = api.ALLREDUCE_FN(_allreduce)
        self._fork()
        out = api.cg_solve_device(self.handle, b.data_ptr(), x.data_ptr(), lam, tol, max_iter, hook)
        self._join()
        return out
