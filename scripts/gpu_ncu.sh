# ncu evidence for profiles/: launch list of one short bench run + full capture of the top kernels
set -x; mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'dgemm_|crossprod_tc' -c 6 -o gpurun_out/prof_top $CMD > gpurun_out/ncu_f.log 2>&1
tail -3 gpurun_out/ncu_l.log gpurun_out/ncu_f.log; ls -la gpurun_out
