#!/bin/bash
# compute-sanitizer over one pass through every kernel family.  usage: sanitize.sh [memcheck|racecheck|synccheck|initcheck]
mkdir -p gpurun_out
TOOL=${1:-memcheck}
python scripts/sanitize_target.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -n 20 gpurun_out/sanitize_plain.log; exit 1; }
timeout 900 compute-sanitizer --tool $TOOL --print-limit 20 --error-exitcode 7 python scripts/sanitize_target.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "sanitizer($TOOL) rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|SANITIZE_TARGET_DONE|Invalid|hazard" gpurun_out/sanitize_$TOOL.log | head -20
