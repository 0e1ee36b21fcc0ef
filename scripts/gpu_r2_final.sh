# end-of-round check on one B200: complete GPU suite + smoke + default bench (own arm)
set -x; mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -m gpu -x -q --timeout 900 ) 2>&1 | tail -10 > gpurun_out/r2_pytest_gpu.txt; cat gpurun_out/r2_pytest_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.txt 2>&1; tail -2 gpurun_out/r2_smoke.txt
timeout 900 python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err; echo rc=$?; tail -3 gpurun_out/r2_bench_final.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench_final.json'))
e=d['e2e']
print('value', d['value'], 'ms/step', d['ms_per_step'], 'frac', d['roofline']['frac'], d['clocks'])
print('e2e', e['ms_per_step'], 'pageable', e['pageable']['ms_per_step'], e['parity']['ok'], e['host_link'])
print('small', d['small']['ms_median'], d['small']['tensor_core_path'], d['small']['ok'])
print('pcg', d['pcg']['ms_per_iteration'], d['pcg']['roofline']['frac'], d['pcg']['e2e']['ms_per_iteration'])
g=d['grm']; print('grm', g['ms'], g['roofline']['frac'], g['e2e']['ms'], g['e2e'].get('pageable'))
PY
