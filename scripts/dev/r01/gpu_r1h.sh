#!/bin/bash
# session H (1 GPU): tb2m fix-up form on C3, ncu of C4 (27 coefficient fields)
mkdir -p gpurun_out
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%r["achieved"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "launches", d["gpu_launches"], (r.get("single_step_kernel") or {}).get("value"))
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
: > gpurun_out/variants.err
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "temporal" > gpurun_out/pytest_tb.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_tb.log
for v in tb2m:128x14f tb2m:128x14; do
  f=gpurun_out/bench_c3_${v/:/_}.json
  timeout 400 python bench.py --steps 100 --warmup 6 --no-cpu --e2e-steps 0 --variant $v > $f 2>> gpurun_out/variants.err; summ "c3 $v" $f
done
for w in c5 c5f32 c2 c1; do
  f=gpurun_out/bench_${w}_auto.json
  timeout 400 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu --e2e-steps 2 > $f 2>> gpurun_out/variants.err; summ "$w auto" $f
done
tail -5 gpurun_out/variants.err
CMD="python bench.py --workload c4 --steps 4 --warmup 3 --no-cpu --e2e-steps 0"
timeout 600 $CMD > gpurun_out/plain_c4.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tma -s 2 -c 1 -o gpurun_out/prof_c4_tma -f $CMD > gpurun_out/ncu_full_c4.log 2>&1
echo "ncu c4 exit $?"; tail -2 gpurun_out/ncu_full_c4.log
