# Power / clock attribution of the dgemm main kernel: nvidia-smi sampled every 50 ms over 300
# back-to-back steps, for the full kernel and (trace build) with the MMAs or the TMEM stores
# skipped.  usage: gpu_power.sh
mkdir -p gpurun_out
run() {  # name, env...
  name=$1; shift
  nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.sw_thermal_slowdown,temperature.gpu --format=csv,noheader,nounits -lms 50 > gpurun_out/power_$name.csv &
  SMI=$!
  env "$@" python bench.py --steps 300 --warmup 3 --no-cpu --no-grm --no-pcg > gpurun_out/pw_$name.json 2>gpurun_out/pw_$name.err
  kill $SMI
  python - $name <<'PY'
import json,sys,statistics as st
n=sys.argv[1]
d=json.load(open('gpurun_out/pw_%s.json'%n))
rows=[l.strip().split(', ') for l in open('gpurun_out/power_%s.csv'%n) if l.strip()]
busy=[r for r in rows if float(r[1])>400]
clk=[float(r[0]) for r in busy]; pw=[float(r[1]) for r in busy]
cap=sum(1 for r in busy if r[2].strip()=='Active')
print("%-10s n_ms %.3f t_ms %.3f frac %.3f | under load: %d samples, SM clock median %.0f MHz (min %.0f max %.0f), power median %.0f W (max %.0f), sw_power_cap active in %d, temp %s C" % (
  n, d['roofline']['n_ms'], d['roofline']['t_ms'], d['roofline']['frac'], len(busy), st.median(clk) if clk else 0, min(clk) if clk else 0, max(clk) if clk else 0, st.median(pw) if pw else 0, max(pw) if pw else 0, cap, busy[-1][4] if busy else '?'))
PY
}
nvidia-smi --query-gpu=power.limit,power.max_limit,clocks.max.sm --format=csv
run full MB200_X=0
run trace MB200_TRACE=1
run nomma MB200_TRACE=1 MB200_DBG=2
run nost MB200_TRACE=1 MB200_DBG=1
run none MB200_TRACE=1 MB200_DBG=3
run s4 MB200_DGEMM_SLICES=4
run s8 MB200_DGEMM_SLICES=8
