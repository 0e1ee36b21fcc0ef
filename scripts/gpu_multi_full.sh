# multi-GPU bench incl. configs[4]: usage  gpurun --gpus N -- 'bash scripts/gpu_multi_full.sh N'
N=${1:-2}; set -x; mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu --grm-full 1 > gpurun_out/bench_full_n$N.json 2> gpurun_out/bench_full_n$N.err
cat gpurun_out/bench_full_n$N.json; tail -8 gpurun_out/bench_full_n$N.err
