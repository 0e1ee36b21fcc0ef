# 8-GPU bench record of the end-of-round state (gpurun --gpus 8): what the driver's scaling run does at N = 8
set -x; mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus 8 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err; echo rc=$?
tail -5 gpurun_out/r2_bench_n8.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench_n8.json').read().strip().splitlines()[-1])
e=d['e2e']
print('value', d['value'], 'ms/step', d['ms_per_step'], 'frac', d['roofline']['frac'], d['clocks'])
print('e2e', e['ms_per_step'], 'pageable', e['pageable']['ms_per_step'], e['parity']['ok'], e['host_link'])
print('pcg', d['pcg']['ms_per_iteration'], d['pcg']['roofline']['frac'], d['pcg']['e2e']['ms_per_iteration'])
for k in ('grm','grm_full'):
    if k in d: print(k, d[k]['ms'], d[k]['roofline']['frac'], json.dumps(d[k].get('e2e'))[:600])
PY
