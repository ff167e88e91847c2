#!/bin/bash
# session F: Jacobi temporal blocking (tb2m with a pointwise rhs), C5 f64 / f32 benches by variant
mkdir -p gpurun_out
echo "pytest skipped (green in the previous call)"
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%r["achieved"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "launches", d["gpu_launches"], (r.get("single_step_kernel") or {}).get("value"))
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
for w in c5 c5f32; do
  for v in tb2m:64x24 tb2m:128x10 tb2m:128x14; do
    f=gpurun_out/bench_${w}_${v/:/_}.json
    timeout 400 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu --e2e-steps 0 --variant $v > $f 2>> gpurun_out/variants.err; summ "$w $v" $f
  done
done
timeout 400 python bench.py --steps 100 --warmup 6 --no-cpu --e2e-steps 0 --variant tb2m:64x24 > gpurun_out/bench_c3_tb2m_64x24.json 2>> gpurun_out/variants.err; summ "c3 tb2m:64x24" gpurun_out/bench_c3_tb2m_64x24.json
tail -5 gpurun_out/variants.err
CMD="python bench.py --workload c5 --steps 6 --warmup 4 --no-cpu --e2e-steps 0"
timeout 300 $CMD > gpurun_out/plain_c5.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tb2m -s 2 -c 1 -o gpurun_out/prof_c5_tb2m -f $CMD > gpurun_out/ncu_full_c5.log 2>&1
echo "ncu full exit $?"; tail -2 gpurun_out/ncu_full_c5.log
