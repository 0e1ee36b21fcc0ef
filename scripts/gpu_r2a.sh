set -x; mkdir -p gpurun_out
bash scripts/dbg/box_facts.sh > gpurun_out/box_facts.txt 2>&1
timeout 300 python scripts/dbg/pageable_e2e.py > gpurun_out/pageable_before.txt 2>&1
STEPS=300 timeout 900 bash scripts/probes/sol_probe.sh > gpurun_out/sol_probe.txt 2>&1
cat gpurun_out/box_facts.txt gpurun_out/pageable_before.txt gpurun_out/sol_probe.txt
