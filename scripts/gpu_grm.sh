set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_crossprod.py -m gpu -x -q 2>&1 | tail -5
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-pcg > gpurun_out/bench_grm.json 2> gpurun_out/bench_grm.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_grm.json'))
print(json.dumps(d['grm'], indent=1))
PY
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_grm.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-pcg > gpurun_out/ncu_grm.log 2>&1
python scripts/ncu_summary.py launches gpurun_out/launches_grm.csv 2>&1 | head -30
