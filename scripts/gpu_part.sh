set -x; mkdir -p gpurun_out; export PYTHONPATH=$PWD
timeout 900 python -m pytest tests/test_gpu_crossprod.py -m gpu -q -x 2>&1 | tail -6 | tee gpurun_out/pytest_part.log
timeout 300 python scripts/dbg/part_check.py 50000 10000 2 1 2>&1 | tee gpurun_out/part_check.log
timeout 300 python scripts/dbg/part_check.py 1000000 100000 8 3 2>&1 | tee -a gpurun_out/part_check.log
