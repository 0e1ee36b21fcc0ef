set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_dgemm.py -m gpu -x -q -k "pipeline or mixed or adaptive" 2>&1 | tail -3
for ch in 4 6; do
MB200_PIPE_CHUNKS=$ch timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-grm --no-pcg > gpurun_out/bench_ch$ch.json 2> gpurun_out/bench_ch$ch.err
python - $ch <<'PY'
import json,sys
d=json.load(open('gpurun_out/bench_ch%s.json' % sys.argv[1]))
print("chunks=%s value %.3e  ms/step %.2f  frac %.3f  e2e %.3e e2e_ms %.2f clk %s" % (sys.argv[1], d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['e2e']['ms_per_step'], d['clocks']['sm_mhz']))
PY
done
