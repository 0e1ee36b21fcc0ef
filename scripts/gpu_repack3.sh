set -x; mkdir -p gpurun_out; export PYTHONPATH=$PWD
timeout 600 python -m pytest tests/test_gpu_dgemm.py tests/test_bedfile.py tests/test_gpu_crossprod.py -m gpu -q -k "repack or bed or full_size or simulator or grm or counts or plink_format" 2>&1 | tail -6 | tee gpurun_out/pytest_repack.log
timeout 300 python scripts/dbg/handle_time.py 3 2>&1 | tee gpurun_out/handle_time.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:repack_fast -c 1 -o gpurun_out/prof_repack_v3 python scripts/dbg/handle_time.py 1 > gpurun_out/ncu_repack.log 2>&1
ncu -i gpurun_out/prof_repack_v3.ncu-rep --page details | grep -E "Duration|DRAM Throughput|Issue Slots Busy|Memory Throughput" 
timeout 600 python bench.py --steps 3 --warmup 3 --no-pcg --no-cpu > gpurun_out/bench_grmlegs.json 2> gpurun_out/bench_grmlegs.err; python -c "
import json; d=json.load(open('gpurun_out/bench_grmlegs.json')); g=d['grm']; print(json.dumps({k:g.get(k) for k in ['value','ms','expand_gemm_ms','epilogue_ms','e2e']}, indent=1)); print(g['roofline']['frac'], d['config'].get('setup_s'))"
