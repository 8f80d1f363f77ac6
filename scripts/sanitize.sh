#!/usr/bin/env bash
# compute-sanitizer target.  Usage: scripts/sanitize.sh [memcheck|racecheck|synccheck|initcheck]
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 120 python scripts/sanitize_target.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/sanitize_plain.log; exit 1; }
cat gpurun_out/sanitize_plain.log | tail -1
timeout ${SANITIZE_TIMEOUT:-200} compute-sanitizer --tool $TOOL --error-exitcode 9 --print-limit 20 python scripts/sanitize_target.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "sanitizer($TOOL) rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|\"ok\"" gpurun_out/sanitize_$TOOL.log | tail -5
grep -E "Invalid|hazard|Barrier error" gpurun_out/sanitize_$TOOL.log | head -10
