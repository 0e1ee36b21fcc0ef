set -x; mkdir -p gpurun_out; export PYTHONPATH=$PWD PART_MODES=2
for rep in 1 2; do
echo "== default"; timeout 300 python scripts/dbg/part_check.py 1000000 100000 8 3 2>&1 | grep mode
echo "== no overlap"; MB200_NO_ROWSUM_OVERLAP=1 timeout 300 python scripts/dbg/part_check.py 1000000 100000 8 3 2>&1 | grep mode
echo "== no overlap, byte expand"; MB200_EXPAND_BYTES=1 MB200_NO_ROWSUM_OVERLAP=1 timeout 300 python scripts/dbg/part_check.py 1000000 100000 8 3 2>&1 | grep mode
echo "== overlap, byte expand"; MB200_EXPAND_BYTES=1 timeout 300 python scripts/dbg/part_check.py 1000000 100000 8 3 2>&1 | grep mode
done 2>&1 | tee gpurun_out/ab_part.log
