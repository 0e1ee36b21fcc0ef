set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dgemm.py tests/test_gpu_solver.py -x -q --timeout 600 2>&1 | tail -5
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q --timeout 300 -k "dgemm or adaptive or pageable or nonfinite or cg" 2>&1 | tail -5
# host-side phases of the sharded handle with 8 shards on one GPU (enqueue costs of 8 threads)
PRINT_LEVEL=1 timeout 600 python bench.py --steps 3 --warmup 3 --no-pcg --no-small --no-cpu --no-grm --e2e-devices 0,0,0,0,0,0,0,0 > gpurun_out/r2g_bench8.json 2> gpurun_out/r2g_bench8.err; echo rc=$?
grep -a "'N'\|'T'" gpurun_out/r2g_bench8.err | tail -24
python -c "
import json; d=json.load(open('gpurun_out/r2g_bench8.json')); print(json.dumps(d['e2e'])[:900]); print(d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms'], d['ms_per_step'])"
timeout 600 python bench.py --steps 10 --warmup 3 --no-pcg --no-small --no-cpu --no-grm 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('default: frac', d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms'], 'step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'pageable', d['e2e']['pageable']['ms_per_step'])"
