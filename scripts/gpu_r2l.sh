set -x; mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_multi.py tests/test_gpu_sharded_nccl.py -x -q --timeout 300 -k "tiled or grm or env or cuda_device or nccl or concurrent or simulated or cg" 2>&1 | tail -8 > gpurun_out/r2l_multi_2gpu.txt; cat gpurun_out/r2l_multi_2gpu.txt
# host-side phases of the sharded handle on two real GPUs (one process, no torchrun)
PRINT_LEVEL=1 timeout 600 python bench.py --steps 3 --warmup 3 --no-pcg --no-small --no-cpu --no-grm --e2e-devices 0,1 > gpurun_out/r2l_phase.json 2> gpurun_out/r2l_phase.err; echo rc=$?
grep -a "'N'\|'T'" gpurun_out/r2l_phase.err | tail -12
python -c "
import json; d=json.load(open('gpurun_out/r2l_phase.json')); e=d['e2e']; print('e2e', e['ms_per_step'], e['n_ms'], e['t_ms'], 'pageable', e['pageable']['ms_per_step'], e['parity'])"
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2l_bench_n2.json 2> gpurun_out/r2l_bench_n2.err; echo rc=$?; tail -3 gpurun_out/r2l_bench_n2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2l_bench_n2.json'))
e=d['e2e']
print('value', d['value'], 'ms/step', d['ms_per_step'], 'frac', d['roofline']['frac'])
print('e2e', e['ms_per_step'], 'n', e['n_ms'], 't', e['t_ms'], 'pageable', e['pageable']['ms_per_step'], e['parity'])
print('pcg', d['pcg']['ms_per_iteration'], d['pcg']['e2e'])
print('grm', d['grm']['ms'], json.dumps(d['grm']['e2e'])[:700])
PY
