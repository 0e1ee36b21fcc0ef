set -x; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_crossprod.py -x -q --timeout 300 -k "tiled or grm or crossprod or counts or ld" 2>&1 | tail -15 > gpurun_out/r2d_tests.txt; cat gpurun_out/r2d_tests.txt
timeout 600 python bench.py --steps 3 --warmup 3 --no-pcg --no-small --no-cpu > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo rc=$?; tail -3 gpurun_out/r2d_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2d_bench.json'))
print(json.dumps(d['grm']['e2e'])[:1200])
PY
export MB200_DGEMM_SLICES=6
for v in "full MB200_X=0" "lowbits MB200_PROBE=8" "full2 MB200_X=0" "lowbits2 MB200_PROBE=8"; do set -- $v; env $2 python bench.py --steps 300 --warmup 3 --no-cpu --no-grm --no-pcg --no-small 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); r=d['roofline']; c=d['clocks']['device_timed']; print('$1', 'n_ms %.3f t_ms %.3f frac %.3f clk %s W %s parity %s' % (r['n_ms'], r['t_ms'], r['frac'], c['sm_mhz'], c['power_w'], d['e2e']['parity']['ok']))"; done
