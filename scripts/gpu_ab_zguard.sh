# same-box A/B: z = generic loads (old), y = explicit ld.shared; guard 0 none (racy), 1 proxy fence, 2 data dependency
for round in 1 2 3; do for v in z0 y0 y1 y2; do
MB200_LIB_PATH=ab/lib_$v.so timeout 600 python bench.py --steps 20 --warmup 3 --no-pcg --no-small --no-cpu --no-grm 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$v round $round: frac %.4f n_ms %.3f t_ms %.3f step %.3f e2e %.3f parity %s' % (d['roofline']['frac'], d['roofline']['n_ms'], d['roofline']['t_ms'], d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['parity']['ok']))"
done; done
MB200_LIB_PATH=ab/lib_y2.so timeout 600 python scripts/dbg/concurrent_check.py 125000 100000 10 4 NT 12 2>&1 | tail -3
