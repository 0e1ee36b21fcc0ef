# launch list of one short bench run on the end-of-round code (per-kernel shares; cold-cache, serialised)
set -x; mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu --no-pcg --e2e-devices 0,0 --wide-ncol 0"
$CMD > gpurun_out/r2_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r2_launches_final.csv $CMD > gpurun_out/r2_ncu_l.log 2>&1
tail -2 gpurun_out/r2_ncu_l.log; wc -l gpurun_out/r2_launches_final.csv
