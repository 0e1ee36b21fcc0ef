set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_solver.py -x -q --timeout 600 -k "full_size" 2>&1 | tail -15
timeout 1200 python -m pytest tests/test_gpu_crossprod.py -x -q --timeout 900 -k "full_size_sampled or out_of_core" 2>&1 | tail -15
CMD="python scripts/dbg/wide_once.py"
$CMD > gpurun_out/r2q_wide_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum --clock-control none -k regex:dgemm_tc_kernel --csv --log-file gpurun_out/r2q_wide_ncu.csv $CMD > gpurun_out/r2q_wide_ncu.log 2>&1
cat gpurun_out/r2q_wide_plain.log | tail -3; grep -a "dgemm_tc" gpurun_out/r2q_wide_ncu.csv | cut -d, -f5,12- | head -20
