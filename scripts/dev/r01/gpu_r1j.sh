#!/bin/bash
# session J (1 GPU): rank-2 temporal blocking (sweep_rows2)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "temporal or row_marching or every_variant" > gpurun_out/pytest_tb2r.log 2>&1; echo "pytest exit $?"; tail -8 gpurun_out/pytest_tb2r.log
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%r["achieved"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "launches", d["gpu_launches"], (r.get("single_step_kernel") or {}).get("value"))
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
: > gpurun_out/variants.err
for w in c2 c1; do
  for v in tb2r:256x3 tb2r:256x6 tb2r:512x3; do
    f=gpurun_out/bench_${w}_${v/:/_}.json
    timeout 300 python bench.py --workload $w --steps 40 --warmup 6 --no-cpu --e2e-steps 0 --variant $v > $f 2>> gpurun_out/variants.err; summ "$w $v" $f
  done
done
tail -5 gpurun_out/variants.err
CMD="python bench.py --workload c2 --steps 6 --warmup 4 --no-cpu --e2e-steps 0"
timeout 300 $CMD > gpurun_out/plain_c2.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sweep_rows2 -s 2 -c 1 -o gpurun_out/prof_c2_rows2 -f $CMD > gpurun_out/ncu_full_c2r2.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/ncu_full_c2r2.log
