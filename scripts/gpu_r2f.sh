set -x; mkdir -p gpurun_out
bash scripts/dbg/box_facts.sh > gpurun_out/box_facts_2gpu.txt 2>&1; cat gpurun_out/box_facts_2gpu.txt
timeout 1500 python -m pytest tests/test_gpu_multi.py -x -q --timeout 300 2>&1 | tail -8 > gpurun_out/r2f_multi_2gpu.txt; cat gpurun_out/r2f_multi_2gpu.txt
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2f_bench_n2.json 2> gpurun_out/r2f_bench_n2.err; echo rc=$?; tail -5 gpurun_out/r2f_bench_n2.err; cat gpurun_out/r2f_bench_n2.json
