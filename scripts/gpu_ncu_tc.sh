set -x; mkdir -p gpurun_out
export MB200_DGEMM_SLICES=${S:-6}
MB200_TRACE=1 timeout 100 python scripts/trace_tc.py 2>&1 | tail -16
CMD="python bench.py --steps 2 --warmup 1 --no-cpu --no-grm"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:dgemm_tc -c 2 -o gpurun_out/prof_tc3 $CMD > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out | tail -3
