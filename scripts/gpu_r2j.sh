set -x
timeout 600 python scripts/dbg/concurrent_check.py 125000 100000 10 4 NT 8 2>&1 | tail -14
MB200_TC_FLAGS=1 timeout 600 python scripts/dbg/concurrent_check.py 125000 100000 10 4 NT 8 2>&1 | tail -6
timeout 600 python scripts/dbg/concurrent_check.py 125000 100000 10 2 N 12 2>&1 | tail -6
timeout 600 python scripts/dbg/concurrent_check.py 125000 100000 10 4 T 8 2>&1 | tail -6
MB200_DGEMM_SLICES=8 timeout 600 python scripts/dbg/concurrent_check.py 125000 100000 10 4 N 8 2>&1 | tail -6
timeout 600 python scripts/dbg/concurrent_check.py 125000 100000 1 4 N 8 2>&1 | tail -6
