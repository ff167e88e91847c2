#!/bin/bash
# session P (1 GPU): small runs of every kernel family (compute-sanitizer is closed on this pool, so this is a plain run)
mkdir -p gpurun_out
timeout 120 python scripts/dev/sanitize_small.py > gpurun_out/sanitize_plain.log 2>&1; echo "plain exit $?"; tail -3 gpurun_out/sanitize_plain.log
