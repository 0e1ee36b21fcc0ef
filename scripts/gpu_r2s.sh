set -x; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dgemm.py -x -q --timeout 300 -k "wide_b or matches_oracle" 2>&1 | tail -12
CMD="python scripts/dbg/wide_once.py"
timeout 300 $CMD > gpurun_out/r2s_wide_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -k regex:dgemm_tc_kernel --csv --log-file gpurun_out/r2s_wide_ncu.csv $CMD > gpurun_out/r2s_wide_ncu.log 2>&1
tail -4 gpurun_out/r2s_wide_plain.log; grep -a "dgemm_tc" gpurun_out/r2s_wide_ncu.csv | cut -d, -f5,12- | head -20
MB200_WIDE_CLUSTER=0 timeout 300 python scripts/dbg/wide_once.py 2>&1 | tail -2
