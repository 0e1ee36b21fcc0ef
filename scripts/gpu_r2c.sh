set -x; mkdir -p gpurun_out
# small-size sanity of every bench leg, single-process e2e arm forced onto two shards of GPU 0
timeout 600 python bench.py --snps 200000 --indiv 20000 --steps 3 --warmup 3 --pcg-snps 100000 --pcg-indiv 20000 --pcg-iters 5 --e2e-devices 0,0 --no-cpu > gpurun_out/r2c_bench_small.json 2> gpurun_out/r2c_bench_small.err; echo rc=$?; tail -5 gpurun_out/r2c_bench_small.err; cat gpurun_out/r2c_bench_small.json
# the default bench line
timeout 900 python bench.py > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo rc=$?; tail -5 gpurun_out/r2c_bench.err; cat gpurun_out/r2c_bench.json
STEPS=400 timeout 1200 bash scripts/probes/sol_probe.sh > gpurun_out/sol_probe.txt 2>&1
cat gpurun_out/sol_probe.txt
