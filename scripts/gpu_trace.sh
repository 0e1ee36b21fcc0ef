set -x; mkdir -p gpurun_out
MB200_TRACE=1 timeout 600 python scripts/trace_tc.py 2>&1 | tail -20
MB200_DGEMM_SLICES=4 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-grm --no-pcg 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('S=4 frac %.3f n_ms %.3f t_ms %.3f' % (d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms']))"
MB200_DGEMM_SLICES=8 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-grm --no-pcg 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('S=8 frac %.3f n_ms %.3f t_ms %.3f' % (d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms']))"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --no-grm --no-pcg --ncol 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('ncol=2 frac %.3f n_ms %.3f t_ms %.3f' % (d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms']))"
