# multi-GPU bench: usage  gpurun --gpus N -- 'bash scripts/gpu_multi.sh N'
N=${1:-2}; set -x; mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
cat gpurun_out/bench_n$N.json; tail -8 gpurun_out/bench_n$N.err
