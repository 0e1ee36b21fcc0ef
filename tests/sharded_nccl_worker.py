"""Worker of tests/test_gpu_sharded_nccl.py (one process per GPU under torch.distributed.run):
ShardedCompressed 'N' / 'T' / CG over NCCL against the oracle, no global synchronisation between the
torch stream and the library stream other than the wrappers' own ordering."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from miraculix_b200 import api  # noqa: E402
from miraculix_b200.sharded import ShardedCompressed  # noqa: E402
from oracle import cg as ocg  # noqa: E402
from oracle import cport, simulate  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    api.set_options(use_gpu=True, not_center=False)
    snps, indiv, ncol = 5003, 1037, 5          # ragged: the last shard is short
    plink, Z, f = simulate.simulate(snps, indiv, seed=21, missing=0.01)
    sh = ShardedCompressed(None, snps, indiv, rank, world)
    b0, cnt = sh.s_begin, sh.s_count
    sh.handle = api.init_compressed(plink[b0:b0 + cnt], cnt, indiv, f[b0:b0 + cnt], ncol)
    rng = np.random.default_rng(3)
    Bn = np.asfortranarray(rng.standard_normal((snps, ncol)))
    Bt = np.asfortranarray(rng.standard_normal((indiv, ncol)))
    # inputs are produced on torch's current stream right before the call (no global sync)
    dBn = torch.from_numpy(np.ascontiguousarray(Bn[b0:b0 + cnt].T)).to(dev, non_blocking=True) * 1.0
    dBt = torch.from_numpy(np.ascontiguousarray(Bt.T)).to(dev, non_blocking=True) * 1.0
    Cn = torch.empty((ncol, indiv), dtype=torch.float64, device=dev)
    Ct = torch.empty((ncol, cnt), dtype=torch.float64, device=dev)
    sh.dgemm_n(dBn, Cn)
    sh.dgemm_t(dBt, Ct)
    got_n = (Cn * 1.0).cpu().numpy().T          # consumed on torch's stream right after
    got_t = (Ct * 1.0).cpu().numpy().T
    want_n = cport.dgemm_compressed("N", plink, snps, indiv, f, Bn, True)
    want_t = cport.dgemm_compressed("T", plink, snps, indiv, f, Bt, True)[b0:b0 + cnt]
    en = np.abs(got_n - want_n).max() / np.abs(want_n).max()
    et = np.abs(got_t - want_t).max() / np.abs(want_t).max()
    assert en <= 1e-10 and et <= 1e-10, (rank, en, et)
    # CG over the shards through the all-reduce hook
    b = torch.from_numpy(rng.standard_normal(indiv)).to(dev)
    x = torch.zeros(indiv, dtype=torch.float64, device=dev)
    lam, iters = 0.05 * snps, 10
    it, res = sh.cg_solve(b, x, lam=lam, tol=0.0, max_iter=iters)
    xo, ito, _ = ocg.cg_solve(plink, snps, indiv, f, b.cpu().numpy(), np.zeros(indiv), lam=lam, tol=0.0,
                              max_iter=iters)
    ex = np.abs(x.cpu().numpy() - xo).max() / np.abs(xo).max()
    assert it == ito == iters and ex <= 1e-8, (rank, it, ex)
    torch.cuda.synchronize()
    dist.barrier()
    sh._stream = None
    api.free_compressed(sh.handle)
    dist.destroy_process_group()
    if rank == 0:
        print(f"sharded nccl ok: N {en:.2e} T {et:.2e} CG {ex:.2e}")


if __name__ == "__main__":
    main()
