"""The one-process-per-GPU arm (miraculix_b200/sharded.py over torch.distributed / NCCL) on real
GPUs: 'N' / 'T' / CG against the oracle at <= 1e-10 / 1e-8 on a ragged shape, two ranks.  NCCL
refuses two ranks on one device, so this needs a two-GPU lease (skipped otherwise); the CPU suite
covers the partition logic with gloo (tests/test_sharded.py), and bench.py checks this arm against
the single-process multi-device arm at every N."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_sharded_products_over_nccl(gpu_lib):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (NCCL: one rank per device)")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29533",
           os.path.join(ROOT, "tests", "sharded_nccl_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "sharded nccl ok" in r.stdout
