set -x
timeout 600 python scripts/dbg/concurrent_check.py 125000 100000 10 4 2>&1 | tail -6
MB200_DGEMM_MODE=0 timeout 900 python scripts/dbg/concurrent_check.py 125000 100000 10 4 2>&1 | tail -6
CUDA_LAUNCH_BLOCKING=1 timeout 600 python scripts/dbg/multi_check.py 1000000 100000 10 8 2>&1 | tail -8
MB200_DGEMM_MODE=0 timeout 900 python scripts/dbg/multi_check.py 400000 100000 10 8 2>&1 | tail -8
