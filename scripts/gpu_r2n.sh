set -x; mkdir -p gpurun_out
bash scripts/dbg/box_facts.sh > gpurun_out/box_facts_8gpu.txt 2>&1; head -12 gpurun_out/box_facts_8gpu.txt
NCCL_DEBUG=WARN timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2n_bench_n8.json 2> gpurun_out/r2n_bench_n8.err; echo rc=$?; tail -5 gpurun_out/r2n_bench_n8.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2n_bench_n8.json'))
e=d['e2e']
print('value', d['value'], 'ms/step', d['ms_per_step'], 'frac', d['roofline']['frac'])
print('e2e', e['ms_per_step'], 'n', e['n_ms'], 't', e['t_ms'], 'pageable', e['pageable']['ms_per_step'], e['parity'])
print('link', e['host_link'], e['host_link_floor_ms_per_step'])
print('pcg', d['pcg']['ms_per_iteration'], d['pcg']['e2e'])
print('grm', d['grm']['ms'], json.dumps(d['grm']['e2e'])[:700])
print('grm_full', d['grm_full']['ms'], d['grm_full']['roofline']['frac'], json.dumps(d['grm_full']['e2e'])[:1200])
PY
