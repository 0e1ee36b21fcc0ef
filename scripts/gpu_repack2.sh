set -x; mkdir -p gpurun_out; export PYTHONPATH=$PWD
timeout 300 python scripts/dbg/handle_time.py 3 2>&1 | tee gpurun_out/handle_time.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:repack_fast -c 1 -o gpurun_out/prof_repack python scripts/dbg/handle_time.py 1 > gpurun_out/ncu_repack.log 2>&1
timeout 300 python -m pytest tests/test_gpu_crossprod.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 3 --warmup 3 --no-pcg --no-cpu > gpurun_out/bench_grmlegs.json 2> gpurun_out/bench_grmlegs.err; python -c "
import json; d=json.load(open('gpurun_out/bench_grmlegs.json')); print(json.dumps({k:d['grm'].get(k) for k in ['e2e','ms']}, indent=1))"
