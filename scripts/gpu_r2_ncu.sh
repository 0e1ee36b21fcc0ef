# ncu evidence for profiles/r2_*: launch list of one short bench run + full capture of the top kernels
set -x; mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu --no-pcg --no-small --e2e-devices 0,0"
$CMD > gpurun_out/r2_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'dgemm_tc_kernel|crossprod_tc_kernel|finish_chunk|reduce_partials|allgather_rows|col_stats' -c 14 -o gpurun_out/r2_prof_top $CMD > gpurun_out/r2_ncu_f.log 2>&1
tail -3 gpurun_out/r2_ncu_l.log gpurun_out/r2_ncu_f.log; ls -la gpurun_out | tail
