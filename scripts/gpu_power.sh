mkdir -p gpurun_out
nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.active --format=csv,noheader -lms 50 > gpurun_out/power.csv &
SMI=$!
MB200_DGEMM_SLICES=${S:-6} python bench.py --steps 300 --warmup 3 --no-cpu --no-grm > gpurun_out/pw.json 2>gpurun_out/pw.err
kill $SMI
python - <<'PY'
import json
d=json.load(open('gpurun_out/pw.json'))
print("value %.3e frac %.3f n_ms %.2f t_ms %.2f" % (d['value'], d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms']))
rows=[l.strip().split(', ') for l in open('gpurun_out/power.csv') if l.strip()]
import collections
print("samples", len(rows))
for r in rows[::8][:60]: print(r)
PY
