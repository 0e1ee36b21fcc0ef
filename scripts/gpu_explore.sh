set -x; mkdir -p gpurun_out
show() { python - "$1" <<'PY'
import json,sys
d=json.load(open(sys.argv[1]))
print(sys.argv[1], "value %.3e  ms/step %.2f  frac %.3f  n_ms %.2f t_ms %.2f" % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms']))
PY
}
MB200_DGEMM_SLICES=6 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-grm > gpurun_out/b_s6.json 2>gpurun_out/b_s6.err; show gpurun_out/b_s6.json
MB200_DGEMM_SLICES=4 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-grm > gpurun_out/b_s4.json 2>gpurun_out/b_s4.err; show gpurun_out/b_s4.json
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-grm --ncol 2 > gpurun_out/b_n2.json 2>gpurun_out/b_n2.err; show gpurun_out/b_n2.json
CMD="python bench.py --steps 2 --warmup 1 --no-cpu --no-grm"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:dgemm_tc -c 2 -o gpurun_out/prof_tc2 $CMD > gpurun_out/ncu_f.log 2>&1
