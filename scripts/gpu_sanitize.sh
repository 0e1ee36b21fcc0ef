# compute-sanitizer, ONE tool per gpurun call (B200_PROFILING.md): TOOL=memcheck|racecheck|synccheck
# on smoke() (plink2compressed + dgemm 'N'/'T' tensor-core path + crossproduct) and one CTA-pair GRM
# partition test and one sharded-handle test.
set -x; mkdir -p gpurun_out
TOOL=${TOOL:-memcheck}
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/san_plain.log 2>&1 &&
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_sanitizer_${TOOL}_smoke.txt 2>&1
tail -12 gpurun_out/r2_sanitizer_${TOOL}_smoke.txt
timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python -m pytest tests/test_gpu_crossprod.py tests/test_gpu_multi.py -x -q --timeout 1200 -k "pairs_match_single_cta and 517 or multi_dgemm_matches_oracle_and_single and 1003 and 2-1-0" > gpurun_out/r2_sanitizer_${TOOL}_tests.txt 2>&1
tail -12 gpurun_out/r2_sanitizer_${TOOL}_tests.txt
