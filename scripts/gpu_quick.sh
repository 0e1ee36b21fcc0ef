set -x; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_dgemm.py -m gpu -x -q 2>&1 | tail -4
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-grm > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print("value %.3e  ms/step %.2f  frac %.3f  n_ms %.2f t_ms %.2f e2e %.3e" % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms'], d['e2e']['value']))
PY
tail -3 gpurun_out/bench_quick.err
