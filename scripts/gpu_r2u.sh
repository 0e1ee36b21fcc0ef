set -x; mkdir -p gpurun_out
timeout 300 python scripts/dbg/chunk_sweep.py 125000 100000 10 2>&1 | tail -4
timeout 300 python scripts/dbg/chunk_sweep.py 1000000 100000 10 2>&1 | tail -4
timeout 900 python -m pytest tests/test_gpu_dgemm.py tests/test_gpu_multi.py tests/test_bedfile.py -x -q --timeout 600 -k "pipeline or pageable or host or mixed or read_bed or multi_dgemm_matches" 2>&1 | tail -4
