# solver tests + one bench with the PCG leg
set -x; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_solver.py -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_solver.log; cat gpurun_out/pytest_solver.log
timeout 900 python bench.py --no-cpu --no-grm --steps 5 > gpurun_out/bench_pcg.json 2> gpurun_out/bench_pcg.err; cat gpurun_out/bench_pcg.json; tail -5 gpurun_out/bench_pcg.err
