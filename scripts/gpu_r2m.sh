set -x; mkdir -p gpurun_out
PRINT_LEVEL=2 timeout 600 python bench.py --steps 2 --warmup 3 --no-pcg --no-small --no-cpu --no-grm --e2e-devices 0,0 > gpurun_out/r2m_phase.json 2> gpurun_out/r2m_phase.err; echo rc=$?
grep -a "pipeline dev\|'N'\|'T'" gpurun_out/r2m_phase.err | tail -60
