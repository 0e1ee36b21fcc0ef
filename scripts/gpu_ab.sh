# A/B two builds of the library on the SAME box: ab/lib_A.so vs ab/lib_B.so
mkdir -p gpurun_out
for rep in 1 2; do for v in A B; do for s in 7 6; do
MB200_LIB_PATH=$PWD/ab/lib_$v.so MB200_DGEMM_SLICES=$s timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-grm > gpurun_out/ab.json 2> gpurun_out/ab.err
python - $v $s <<'PY'
import json,sys
d=json.load(open('gpurun_out/ab.json'))
print("%s S=%s value %.3e frac %.3f n_ms %.2f t_ms %.2f clk %s" % (sys.argv[1], sys.argv[2], d['value'], d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms'], d['clocks']['sm_mhz']))
PY
done; done; done
