# A/B of MB200_TC_FLAGS values on the SAME box: usage gpu_ab_flags.sh "0 2" [extra env]
set -x; mkdir -p gpurun_out
true
for rep in 1 2; do for fl in $1; do
MB200_TC_FLAGS=$fl timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-grm --no-pcg > gpurun_out/ab.json 2> gpurun_out/ab.err
python - $fl <<'PY'
import json,sys
d=json.load(open('gpurun_out/ab.json'))
print("flags=%s value %.3e frac %.3f n_ms %.3f t_ms %.3f e2e %.3e clk %s" % (sys.argv[1], d['value'], d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms'], d['e2e']['value'], d['clocks']['sm_mhz']))
PY
done; done
