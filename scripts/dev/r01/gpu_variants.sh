#!/bin/bash
# usage: gpu_variants.sh <workload> <variant>...   parity tests first, then one bench line per kernel variant
mkdir -p gpurun_out
W=$1; shift
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log
for v in "$@"; do
  for zc in ${ZCHUNKS:-0}; do
  f=gpurun_out/bench_${W}_${v/:/_}_z${zc}.json
  timeout 300 python bench.py --workload $W --steps ${STEPS:-50} --warmup 5 --no-cpu --e2e-steps 1 --variant $v --zchunk $zc > $f 2>> gpurun_out/variants.err
  python - "$v z$zc" $f <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); print(sys.argv[1], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%d["roofline"]["achieved"], "frac %.3f"%d["roofline"]["frac"], "ms %.4f"%d["ms_per_step"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
  done
done
tail -5 gpurun_out/variants.err
