#!/bin/bash
# session I (1 GPU): restructured sweep_rows (steady-state fast path), row-wise late loads for many-input bodies (C4)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/pytest_gpu.log
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
  for v in rows:512 rows:256 rows:1024; do
    f=gpurun_out/bench_${w}_${v/:/_}.json
    timeout 400 python bench.py --workload $w --steps 40 --warmup 5 --no-cpu --e2e-steps 0 --variant $v > $f 2>> gpurun_out/variants.err; summ "$w $v" $f
  done
done
for v in tma:64x8 tma:64x7 tma:64x14; do
  f=gpurun_out/bench_c4_${v/:/_}.json
  timeout 900 python bench.py --workload c4 --steps 12 --warmup 3 --no-cpu --e2e-steps 0 --variant $v > $f 2>> gpurun_out/variants.err; summ "c4 $v" $f
done
tail -5 gpurun_out/variants.err
CMD="python bench.py --workload c2 --steps 6 --warmup 4 --no-cpu --e2e-steps 0"
timeout 300 $CMD > gpurun_out/plain_c2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_rows -s 2 -c 1 -o gpurun_out/prof_c2_rows_v2 -f $CMD > gpurun_out/ncu_full_c2.log 2>&1
echo "ncu c2 exit $?"; tail -2 gpurun_out/ncu_full_c2.log
