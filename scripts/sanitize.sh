#!/bin/bash
# compute-sanitizer pass over CI-sized launches of every kernel family.  One tool per gpurun call, and
# only after the same program has passed without it:   gpurun --timeout 1500 -- 'bash scripts/sanitize.sh memcheck'
# tools: memcheck | racecheck | synccheck | initcheck.  Single GPU only.  Output: gpurun_out/sanitize_<tool>.log
tool=${1:-memcheck}
shift
sections="$@"
mkdir -p gpurun_out
export PYTHONPATH=.
timeout 600 python scripts/sanitize_targets.py $sections > gpurun_out/sanitize_plain.log 2>&1
rc=$?
tail -3 gpurun_out/sanitize_plain.log
if [ $rc -ne 0 ]; then echo "plain run failed rc=$rc"; exit $rc; fi
timeout ${SAN_TIMEOUT:-1500} compute-sanitizer --tool $tool --error-exitcode 99 --launch-timeout 300 \
  python scripts/sanitize_targets.py $sections > gpurun_out/sanitize_$tool.log 2>&1
rc=$?
echo "sanitize $tool rc=$rc"
grep -E "ERROR SUMMARY|Invalid|Race|hazard|section|Uninitialized|SANITIZE_TARGETS_OK" gpurun_out/sanitize_$tool.log 2>/dev/null | tail -30
exit $rc
