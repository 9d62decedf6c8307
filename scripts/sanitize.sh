#!/bin/bash
# compute-sanitizer pass over CI-sized kernel tests.  One tool per gpurun call, and only after the same
# command has passed without it:   gpurun --timeout 900 -- 'bash scripts/sanitize.sh memcheck'
# tools: memcheck | racecheck | synccheck | initcheck.  Single GPU only.  Output: gpurun_out/sanitize_<tool>.log
tool=${1:-memcheck}
sel=${2:-"tests/test_kmeans_gpu.py tests/test_gram_gpu.py tests/test_mf_gpu.py tests/test_lda_gpu.py tests/test_svm_native.py"}
mkdir -p gpurun_out
export HB_SANITIZE=1          # tests may shrink their shapes when this is set
timeout 600 python -m pytest $sel -x -q -m gpu > gpurun_out/sanitize_plain.log 2>&1 && \
timeout 1500 compute-sanitizer --tool $tool --error-exitcode 99 --launch-timeout 120 \
  python -m pytest $sel -x -q -m gpu > gpurun_out/sanitize_$tool.log 2>&1
rc=$?
echo "sanitize $tool rc=$rc"
tail -5 gpurun_out/sanitize_plain.log
grep -E "ERROR SUMMARY|Invalid|Race|hazard|passed|failed" gpurun_out/sanitize_$tool.log 2>/dev/null | tail -12
exit $rc
