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
