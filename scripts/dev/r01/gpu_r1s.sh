#!/bin/bash
# session S (1 GPU): two plane ranges in one launch (neptune_rt_apply_planes2) + regression of the defaults
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log
for w in c3 c2; do
  timeout 300 python bench.py --workload $w --steps 40 --warmup 6 --no-cpu --e2e-steps 0 > gpurun_out/bench_${w}_s.json 2> gpurun_out/bench_${w}_s.err
  python - gpurun_out/bench_${w}_s.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(d["config"]["workload"][:20], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "ms %.4f"%d["ms_per_step"])
except Exception as e: print("failed", e)
PY
done
