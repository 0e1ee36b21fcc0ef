set -x
timeout 600 python scripts/dbg/concurrent_check.py 125000 100000 10 4 NT 12 2>&1 | tail -4
timeout 600 python scripts/dbg/multi_check.py 1000000 100000 10 8 2>&1 | tail -8
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q --timeout 300 -k "concurrent" 2>&1 | tail -3
for i in 1 2; do timeout 600 python bench.py --steps 10 --warmup 3 --no-pcg --no-small --no-cpu --no-grm 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('default: frac', d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms'], 'step', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'pageable', d['e2e']['pageable']['ms_per_step'], d['e2e']['parity']['ok'])"; done
