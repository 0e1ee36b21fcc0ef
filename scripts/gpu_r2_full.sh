# complete GPU suite + smoke + default bench on one B200 (what the driver runs at round end)
set -x; mkdir -p gpurun_out
( time timeout 2400 python -m pytest tests -m gpu -x -q --timeout 900 ) 2>&1 | tail -10 > gpurun_out/r2_pytest_gpu.txt; cat gpurun_out/r2_pytest_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.txt 2>&1; tail -2 gpurun_out/r2_smoke.txt
timeout 900 python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo rc=$?; tail -3 gpurun_out/r2_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2>/dev/null; cut -c1-400 gpurun_out/r2_bench_reference.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_bench.json'))
e=d['e2e']
print('value', d['value'], 'ms/step', d['ms_per_step'], 'frac', d['roofline']['frac'], d['clocks'])
print('e2e', e['ms_per_step'], 'pageable', e['pageable']['ms_per_step'], e['parity']['ok'], e['host_link'])
print('small', d['small']); print('wide', d['wide_b'])
print('pcg', d['pcg']['ms_per_iteration'], d['pcg']['roofline']['frac'], d['pcg']['e2e']['ms_per_iteration'])
print('grm', d['grm']['ms'], d['grm']['roofline']['frac'], json.dumps(d['grm']['e2e'])[:700])
print('cpu', d['cpu_baseline'])
PY
