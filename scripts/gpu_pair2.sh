mkdir -p gpurun_out
for m in 2; do for pf in 0 -6 -7; do
MB200_GRM_PREFETCH=$pf MB200_CROSSPROD_MODE=$m timeout 300 python bench.py --steps 3 --warmup 1 --no-cpu --no-pcg --snps 100000 --indiv 20000 > gpurun_out/pair_$m.json 2> gpurun_out/pair_$m.err || tail -5 gpurun_out/pair_$m.err
python - $m $pf <<'PY'
import json,sys
d=json.load(open('gpurun_out/pair_%s.json'%sys.argv[1]))['grm']
print("mode=%s prefetch=%s GRM ms %.3f (expand+gemm %.3f, rowsums %.3f)  TOPS %.0f  frac(executed) %.3f" % (sys.argv[1], sys.argv[2], d['ms'], d['expand_gemm_ms'], d['epilogue_ms'], d['value'], d['roofline']['frac']))
PY
done; done
