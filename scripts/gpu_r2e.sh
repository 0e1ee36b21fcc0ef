set -x; mkdir -p gpurun_out
timeout 300 python scripts/dbg/grm_e2e_phases.py 50000 10000 1 > gpurun_out/r2e_phases_pinned.txt 2>&1; cat gpurun_out/r2e_phases_pinned.txt | tail -30
timeout 300 python scripts/dbg/grm_e2e_phases.py 50000 10000 0 > gpurun_out/r2e_phases_pageable.txt 2>&1; cat gpurun_out/r2e_phases_pageable.txt | tail -12
timeout 600 python -m pytest tests/test_gpu_multi.py -x -q --timeout 300 -k "tiled or grm" 2>&1 | tail -5
