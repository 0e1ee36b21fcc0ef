set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dgemm.py -x -q --timeout 600 2>&1 | tail -4
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_solver.py -x -q --timeout 600 -k "dgemm or cg or concurrent" 2>&1 | tail -4
timeout 900 python bench.py --no-grm --no-pcg --no-cpu > gpurun_out/r2p_bench.json 2> gpurun_out/r2p_bench.err; echo rc=$?; tail -3 gpurun_out/r2p_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2p_bench.json'))
print('frac', d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms'], 'step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['e2e']['parity']['ok'])
print('wide', json.dumps(d.get('wide_b')))
PY
