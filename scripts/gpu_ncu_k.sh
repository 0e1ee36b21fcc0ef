# usage: bash scripts/gpu_ncu_k.sh <kernel-regex> <outfile-stem> [bench args]
K=$1; O=$2; shift 2
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:$K -c 2 -o gpurun_out/$O python bench.py --steps 2 --warmup 1 --no-cpu --no-pcg "$@" > gpurun_out/$O.log 2>&1
ls -la gpurun_out/$O.ncu-rep
