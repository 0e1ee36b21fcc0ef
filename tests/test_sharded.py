"""CPU tests of the SNP-block sharding host logic (miraculix_b200/sharded.py): partition rules and
a world-size-2 gloo run of both directions with the oracle injected as the rank-local product
(the library itself has no CPU path).  SURVEY.md 8(e)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_rules():
    from miraculix_b200.sharded import BLOCK, partition_snps
    for snps, world in [(1_000_000, 8), (1_000_000, 1), (10_000, 4), (513, 2), (512, 4), (5, 8), (0, 3)]:
        parts = partition_snps(snps, world)
        assert len(parts) == world
        assert sum(c for _, c in parts) == snps
        pos = 0
        for b, c in parts:
            assert b == pos and c >= 0
            if b + c < snps:
                assert (b + c) % BLOCK == 0  # interior boundaries on 128-byte lines
            pos += c
        counts = [c for _, c in parts]
        assert max(counts) - min(counts) < 2 * BLOCK  # balanced to one block (+ the ragged tail)
    with pytest.raises(ValueError):
        partition_snps(10, 0)


def _worker(rank, world, port, snps, indiv, ncol, out):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from miraculix_b200.sharded import ShardedCompressed
    from oracle import cport, simulate
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    plink, Z, f = simulate.simulate(snps, indiv, seed=3, missing=0.01)
    sh = ShardedCompressed(None, snps, indiv, rank, world)
    b, c = sh.s_begin, sh.s_count
    plink_loc, f_loc = plink[b:b + c], f[b:b + c]

    def local(trans, Bt, Ct):  # panels are (n, rows) row-major = column-major rows x n
        B = np.asfortranarray(Bt.numpy().T)
        res = cport.dgemm_compressed("T" if trans else "N", plink_loc, c, indiv, f_loc, B, True)
        Ct.copy_(torch.from_numpy(np.ascontiguousarray(res.T)))

    sh._local = local
    rng = np.random.default_rng(5)
    Bn = rng.standard_normal((snps, ncol))
    Bt = rng.standard_normal((indiv, ncol))
    Cn = torch.zeros((ncol, indiv), dtype=torch.float64)
    sh.dgemm_n(torch.from_numpy(np.ascontiguousarray(Bn[b:b + c].T)), Cn)
    Ct = torch.zeros((ncol, c), dtype=torch.float64)
    sh.dgemm_t(torch.from_numpy(np.ascontiguousarray(Bt.T)), Ct)
    gathered = [None] * world
    dist.all_gather_object(gathered, Ct.numpy().T.copy())
    if rank == 0:
        full_n = cport.dgemm_compressed("N", plink, snps, indiv, f, np.asfortranarray(Bn), True)
        full_t = cport.dgemm_compressed("T", plink, snps, indiv, f, np.asfortranarray(Bt), True)
        en = np.abs(Cn.numpy().T - full_n).max() / np.abs(full_n).max()
        et = np.abs(np.concatenate(gathered, axis=0) - full_t).max() / np.abs(full_t).max()
        out.put((en, et))
    dist.destroy_process_group()


def test_world2_gloo_matches_unsharded():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, 1300, 97, 3, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    en, et = out.get(timeout=10)
    assert en <= 1e-12 and et <= 1e-12, (en, et)


def test_grm_tile_partition_rules():
    """Host rule of the 2-D output-tile partition: every row owned exactly once, row counts agree
    with the library's mb200_crossprod_partition_rows (a host-only query, no GPU needed)."""
    from miraculix_b200 import api
    from miraculix_b200.sharded import TILE, partition_tile_rows
    lib = api.load_library()
    for rows, nparts in [(100_000, 8), (10_000, 1), (10_000, 3), (700, 3), (128, 2), (1, 4), (257, 8)]:
        owned = []
        for part in range(nparts):
            idx = partition_tile_rows(rows, part, nparts)
            assert idx.size == lib.mb200_crossprod_partition_rows(rows, part, nparts)
            assert idx.size % TILE == 0
            assert (np.diff(idx) > 0).all()
            owned.append(idx[idx < rows])
        owned = np.sort(np.concatenate(owned))
        assert np.array_equal(owned, np.arange(rows))
        sizes = [partition_tile_rows(rows, p, nparts).size for p in range(nparts)]
        assert max(sizes) - min(sizes) <= TILE  # balanced to one tile
    with pytest.raises(ValueError):
        partition_tile_rows(10, 2, 2)


def _grm_worker(rank, world, port, snps, indiv, out):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from miraculix_b200.sharded import assemble_grm_panels, partition_tile_rows
    from oracle import cport, simulate
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    plink, Z, f = simulate.simulate(snps, indiv, seed=11)
    x = simulate.pack_2bit(Z)
    # the oracle as the rank-local product: this rank's rows of the centred GRM, written the way
    # mb200_crossprod_partition_device writes a panel (zeros left of the diagonal 256-tile)
    full = cport.grm_epilogue(cport.crossprod(x, snps, indiv), f, True)
    idx = partition_tile_rows(indiv, rank, world)
    panel = np.zeros((idx.size, indiv))
    for lr, i in enumerate(idx):
        if i < indiv:
            first = (i // 256) * 256
            panel[lr, first:] = full[i, first:]
    gathered = [None] * world
    dist.all_gather_object(gathered, panel)  # the only exchange: collecting the finished panels
    if rank == 0:
        got = assemble_grm_panels(gathered, indiv)
        out.put(bool(np.array_equal(got, full)) and bool(np.array_equal(got, got.T)))
    dist.destroy_process_group()


def test_world2_gloo_grm_panels_assemble():
    """configs[4]'s partition on two ranks: panels assemble to the unsharded GRM bit for bit."""
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_grm_worker, args=(r, 2, port, 301, 700, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert out.get(timeout=10)
