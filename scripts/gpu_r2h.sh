set -x
timeout 600 python scripts/dbg/multi_check.py 1000000 100000 10 8 2>&1 | tail -20
MB200_DGEMM_SLICES=8 timeout 600 python scripts/dbg/multi_check.py 1000000 100000 10 8 2>&1 | tail -12
timeout 600 python scripts/dbg/multi_check.py 1000000 100000 10 4 2>&1 | tail -12
