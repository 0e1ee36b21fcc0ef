# e2e vs device-resident step at the per-rank shard size of the 8-GPU run (125k SNPs x 100k), 1 GPU
mkdir -p gpurun_out
for ch in 1 2 4; do
MB200_PIPE_CHUNKS=$ch timeout 600 python bench.py --snps 125000 --steps 20 --warmup 3 --no-cpu --no-grm --no-pcg > gpurun_out/e2s.json 2> gpurun_out/e2s.err || tail -5 gpurun_out/e2s.err
python - $ch <<'PY'
import json,sys
d=json.load(open('gpurun_out/e2s.json'))
print("chunks=%s device ms/step %.3f  e2e ms/step %.3f  (n_ms %.3f t_ms %.3f)" % (sys.argv[1], d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['n_ms'], d['roofline']['t_ms']))
PY
done
