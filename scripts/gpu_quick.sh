set -x; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dgemm.py -m gpu -x -q 2>&1 | tail -4
for s in 7 6; do
MB200_DGEMM_SLICES=$s timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-grm > gpurun_out/bench_s$s.json 2> gpurun_out/bench_s$s.err
python - $s <<'PY'
import json,sys
d=json.load(open('gpurun_out/bench_s%s.json' % sys.argv[1]))
print("S=%s value %.3e  ms/step %.2f  frac %.3f  n_ms %.2f t_ms %.2f e2e %.3e clk %s" % (sys.argv[1], d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms'], d['e2e']['value'], d['clocks']['sm_mhz']))
PY
tail -3 gpurun_out/bench_s$s.err
done
