set -x; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dgemm.py -x -q --timeout 300 -k "wide_b or matches_oracle" 2>&1 | tail -5
timeout 300 python scripts/dbg/wide_once.py 2>&1 | tail -2
MB200_WIDE_CLUSTER=0 timeout 300 python scripts/dbg/wide_once.py 2>&1 | tail -2
