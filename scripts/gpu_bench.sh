set -x; mkdir -p gpurun_out
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
MB200_DGEMM_SLICES=6 timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu --no-grm > gpurun_out/bench_s6.json 2> gpurun_out/bench_s6.err
cat gpurun_out/bench_s6.json; tail -5 gpurun_out/bench_s6.err
