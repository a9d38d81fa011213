#!/bin/bash
# compute-sanitizer over the CI-sized kernel cases; log -> gpurun_out/sanitizer.log (copy to profiles/<round>_sanitizer.log)
set -u
mkdir -p gpurun_out
LOG=gpurun_out/sanitizer.log
: > $LOG
for tool in memcheck synccheck racecheck; do
  echo "==== compute-sanitizer --tool $tool ====" >> $LOG
  timeout ${SAN_TIMEOUT:-600} compute-sanitizer --tool $tool --print-limit 20 --error-exitcode 9 \
      python tools/sanitize_cases.py "$@" 2>&1 | grep -v "^$" | tail -n 60 >> $LOG
  echo "==== exit ${PIPESTATUS[0]} ====" >> $LOG
done
tail -n 40 $LOG
