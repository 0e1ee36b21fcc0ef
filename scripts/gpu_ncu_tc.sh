set -x; mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu --no-grm"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:dgemm_tc -c 2 -o gpurun_out/prof_tc $CMD > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out
