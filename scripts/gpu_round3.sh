# Final-state run on one B200: tests, smoke, bench (both arms), launch list, full ncu capture of the GRM kernels.
set -x; mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/pytest.log; cat gpurun_out/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; cat gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 300 python scripts/dbg/part_check.py 50000 10000 8 3 2>&1 | tee gpurun_out/part_check8.log
CMD="python bench.py --steps 2 --warmup 1 --no-cpu"
timeout 600 $CMD > gpurun_out/ncu_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'crossprod_tc_kernel|expand_2bit' -c 4 -o gpurun_out/prof_grm $CMD --no-pcg --snps 51200 --indiv 10240 > gpurun_out/ncu_f2.log 2>&1
ls -la gpurun_out | tail -8
