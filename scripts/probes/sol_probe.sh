# Speed-of-light split of dgemm_tc_kernel: the SAME kernel (PROBE instantiation,
# no clock64 bookkeeping) with parts of the pipeline switched off, 300 back-to-back steps of
# configs[2] ('n' + 't', ncol = 10) each, nvidia-smi power / SM clock sampled every 50 ms.
#   MB200_PROBE bits: 1 = no tcgen05.st, 2 = no MMAs, 4 = no expansion ALU
# usage (GPU box): bash scripts/probes/sol_probe.sh > gpurun_out/sol_probe.txt
mkdir -p gpurun_out
STEPS=${STEPS:-400}
export MB200_DGEMM_SLICES=${SLICES:-6}   # fixed digits: no adaptive check, no second pass on garbage results
run() {  # name, env...
  name=$1; shift
  env "$@" python bench.py --steps $STEPS --warmup 3 --no-cpu --no-grm --no-pcg --ncol ${NCOL:-10} > gpurun_out/sol_$name.json 2>gpurun_out/sol_$name.err
  python - $name <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open('gpurun_out/sol_%s.json'%n))
except Exception as e:
    print("%-14s FAILED %r" % (n, e)); sys.exit(0)
r=d['roofline']; c=(d.get('clocks') or {}).get('device_timed') or {}
print("%-14s n_ms %.3f t_ms %.3f mean %.3f (x%.2f of the 3.884 ms HBM floor) | timed region: SM clock median %s MHz, power median %s W, reasons %s, %s samples" % (
  n, r['n_ms'], r['t_ms'], (r['n_ms']+r['t_ms'])/2, (r['n_ms']+r['t_ms'])/2/3.884, c.get('sm_mhz'), c.get('power_w'), c.get('reasons'), c.get('samples')))
PY
}
nvidia-smi --query-gpu=name,power.limit,power.max_limit,clocks.max.sm,clocks.max.mem --format=csv
echo "# variant: what still runs"
run full          MB200_X=0                      # product kernel
run probe_all     MB200_PROBE=0                  # PROBE instantiation, everything on (control)
run tma_only      MB200_PROBE=7                  # TMA + LDS + barriers; no ALU, no st, no MMA
run expand_only   MB200_PROBE=2                  # TMA + LDS + ALU + tcgen05.st; no MMA
run mma_only      MB200_PROBE=5                  # TMA + LDS + MMAs on stale TMEM; no ALU, no st
run mma_st        MB200_PROBE=4                  # TMA + LDS + st (raw words) + MMAs; no ALU
run alu_mma       MB200_PROBE=1                  # TMA + LDS + ALU + MMAs; no st
run st_only       MB200_PROBE=6                  # TMA + LDS + st (raw words); no ALU, no MMA
run s8_full       MB200_DGEMM_SLICES=8
run s4_full       MB200_DGEMM_SLICES=4
NCOL=2 run ncol2_full MB200_X=0
run lowbits_full  MB200_PROBE=8                  # everything on, A operand = genotype 0/1/2 (correct results)
