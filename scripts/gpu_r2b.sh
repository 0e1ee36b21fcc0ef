set -x; mkdir -p gpurun_out
nvidia-smi -L
timeout 1200 python -m pytest tests/test_gpu_multi.py -x -q --timeout 240 2>&1 | tail -40 > gpurun_out/r2b_multi.txt; cat gpurun_out/r2b_multi.txt
timeout 900 python -m pytest tests/test_gpu_dgemm.py tests/test_gpu_crossprod.py tests/test_gpu_solver.py -x -q --timeout 600 2>&1 | tail -15 > gpurun_out/r2b_regress.txt; cat gpurun_out/r2b_regress.txt
timeout 300 python scripts/dbg/pageable_e2e.py > gpurun_out/pageable_after.txt 2>&1; cat gpurun_out/pageable_after.txt
