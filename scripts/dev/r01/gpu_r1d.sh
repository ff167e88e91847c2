#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -12 gpurun_out/pytest_gpu.log
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%r["achieved"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "e2e", (d.get("e2e") or {}).get("value"), (d.get("e2e") or {}).get("ms_per_step"), "launches", d["gpu_launches"], d["e2e_gpu_launches"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
timeout 600 python bench.py --no-cpu > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; summ default gpurun_out/bench_default.json; tail -3 gpurun_out/bench_default.err
for v in tb2m:128x14 tb2m:64x16 tb2m:128x14n8; do
  timeout 300 python bench.py --steps 100 --warmup 6 --no-cpu --e2e-steps 0 --variant $v > gpurun_out/bench_c3_${v/:/_}.json 2>> gpurun_out/variants.err; summ $v gpurun_out/bench_c3_${v/:/_}.json
done
tail -3 gpurun_out/variants.err
timeout 900 python bench.py --workload c4 --steps 20 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; summ c4 gpurun_out/bench_c4.json; tail -3 gpurun_out/bench_c4.err
for w in c2 c5; do
  timeout 400 python bench.py --workload $w --steps 30 --warmup 5 --no-cpu --e2e-steps 3 > gpurun_out/bench_${w}.json 2> gpurun_out/bench_${w}.err || tail -3 gpurun_out/bench_${w}.err
  summ $w gpurun_out/bench_${w}.json
done
CMD="python bench.py --steps 8 --warmup 4 --no-cpu --e2e-steps 0 --variant tb2m:128x14"
timeout 300 $CMD > gpurun_out/plain_tb2m.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tb2m -s 2 -c 1 -o gpurun_out/prof_c3_tb2m -f $CMD > gpurun_out/ncu_full_tb2m.log 2>&1
echo "ncu full exit $?"; tail -2 gpurun_out/ncu_full_tb2m.log
