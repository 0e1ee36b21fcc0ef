# configs[4]-style GRM over N GPUs at a reduced size first, then (N == 8) the full bench line
N=${1:-2}; set -x; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_crossprod.py -x -q 2>&1 | tail -3
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 3 --warmup 3 --no-cpu --no-pcg --grm-full 1 --grm-full-snps ${2:-400000} --grm-full-indiv ${3:-40000} > gpurun_out/grmfull_n$N.json 2> gpurun_out/grmfull_n$N.err
tail -5 gpurun_out/grmfull_n$N.err
python - $N <<'PY'
import json,sys
d=json.load(open('gpurun_out/grmfull_n%s.json'%sys.argv[1]))
print(json.dumps(d.get('grm_full'), indent=1)); print(json.dumps(d.get('grm'), indent=1))
PY
