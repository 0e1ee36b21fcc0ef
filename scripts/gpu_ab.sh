# A/B of two prebuilt libraries (ab/lib_A.so, ab/lib_B.so) on the SAME box; env EXTRA passes options
set -x; mkdir -p gpurun_out
for rep in 1 2 3; do for v in A B; do
env $EXTRA MB200_LIB_PATH=$PWD/ab/lib_$v.so timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-grm --no-pcg > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err || tail -5 gpurun_out/ab_$v.err
python - $v <<'PY'
import json,sys
d=json.load(open('gpurun_out/ab_%s.json' % sys.argv[1]))
print("lib=%s value %.3e frac %.3f n_ms %.3f t_ms %.3f e2e %.3e clk %s" % (sys.argv[1], d['value'], d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms'], d['e2e']['value'], d['clocks']['sm_mhz']))
PY
done; done
[ -n "$NOTEST" ] || timeout 900 python -m pytest tests/test_gpu_dgemm.py -x -q 2>&1 | tail -3
