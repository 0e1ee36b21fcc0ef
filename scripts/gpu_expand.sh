set -x; mkdir -p gpurun_out; export PYTHONPATH=$PWD
timeout 900 python -m pytest tests/test_gpu_crossprod.py -m gpu -q -x 2>&1 | tail -6 | tee gpurun_out/pytest_expand.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --steps 3 --warmup 3 --no-pcg --no-cpu > gpurun_out/bench_grmlegs.json 2> gpurun_out/bench_grmlegs.err; python -c "
import json; d=json.load(open('gpurun_out/bench_grmlegs.json')); g=d['grm']; print(json.dumps({k:g.get(k) for k in ['value','ms','expand_gemm_ms','epilogue_ms','e2e']}, indent=1)); print(g['roofline']['frac'])"
timeout 300 python scripts/dbg/part_check.py 1000000 100000 8 3 2>&1 | tee gpurun_out/part_check.log
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:expand_2bit -c 3 python bench.py --steps 1 --warmup 1 --no-pcg --no-cpu --snps 51200 --indiv 10240 2>&1 | grep -E "expand_2bit|duration|dram__" | tee gpurun_out/ncu_expand.log
