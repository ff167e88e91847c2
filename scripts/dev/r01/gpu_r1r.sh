#!/bin/bash
# session R (1 GPU): pipelined rank-2 temporal blocking (tb2r:256x3p)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "temporal or repeat" > gpurun_out/pytest_tb2rp.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_tb2rp.log
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%r["achieved"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "launches", d["gpu_launches"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
: > gpurun_out/variants.err
for w in c2 c1; do
  for v in tb2r:256x3p tb2r:256x3; do
    f=gpurun_out/bench_${w}_${v/:/_}.json
    timeout 300 python bench.py --workload $w --steps 40 --warmup 6 --no-cpu --e2e-steps 0 --variant $v > $f 2>> gpurun_out/variants.err; summ "$w $v" $f
  done
done
tail -3 gpurun_out/variants.err
