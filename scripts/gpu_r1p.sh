#!/bin/bash
# session P (1 GPU): compute-sanitizer passes over small runs of every kernel family
mkdir -p gpurun_out
timeout 120 python scripts/dev/sanitize_small.py > gpurun_out/sanitize_plain.log 2>&1; echo "plain exit $?"; tail -3 gpurun_out/sanitize_plain.log
for tool in memcheck synccheck racecheck; do
  timeout 420 compute-sanitizer --tool $tool --print-limit 20 python scripts/dev/sanitize_small.py > gpurun_out/sanitize_$tool.log 2>&1; echo "$tool exit $?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|bit-exact|MISMATCH|done" gpurun_out/sanitize_$tool.log | sort | uniq -c | head -20
done
