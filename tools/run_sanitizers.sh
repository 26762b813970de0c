#!/bin/bash
# compute-sanitizer over tools/sanitizer_workload.py, one tool after the other; summaries into gpurun_out/ (copied to
# profiles/ by hand). Each tool under its own timeout: racecheck slows shared-memory code by two orders of magnitude.
out=${1:-gpurun_out}
for tool in memcheck synccheck racecheck initcheck; do
  extra=""
  [ "$tool" = racecheck ] && extra="--racecheck-report all"
  [ "$tool" = racecheck ] && wl="--quick" || wl=""
  start=$(date +%s)
  timeout ${SAN_TIMEOUT:-1500} compute-sanitizer --tool $tool $extra --print-limit 20 python tools/sanitizer_workload.py $wl > $out/r02_sanitizer_$tool.log 2>&1
  rc=$?
  echo "== $tool rc=$rc seconds=$(( $(date +%s) - start ))" | tee -a $out/r02_sanitizer_summary.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitizer workload ok|Error:|Race reported|hazard" $out/r02_sanitizer_$tool.log | sort | uniq -c | sort -rn | head -12 | tee -a $out/r02_sanitizer_summary.txt
done
