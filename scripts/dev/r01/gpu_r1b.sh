#!/bin/bash
# re-entry pass: parity, bench lines for every workload (single-step and temporally blocked c3),
# launch list + full ncu capture of the temporally blocked c3 kernel and of the single-step kernel
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > gpurun_out/gpu.txt
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%r["achieved"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "e2e %.2f"%d["e2e"]["value"], "cpu", (d.get("cpu_baseline") or {}).get("value"), d["clocks"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; summ default gpurun_out/bench_default.json
for v in tb2:64x16 tb2:128x16 tb2:64x32 tb2:64x16n7; do
  timeout 300 python bench.py --steps 100 --warmup 6 --no-cpu --e2e-steps 1 --variant $v > gpurun_out/bench_c3_${v/:/_}.json 2>> gpurun_out/variants.err; summ $v gpurun_out/bench_c3_${v/:/_}.json
done
timeout 300 python bench.py --steps 100 --warmup 6 --no-cpu --e2e-steps 1 --tblock 1 > gpurun_out/bench_c3_single.json 2>> gpurun_out/variants.err; summ c3-single gpurun_out/bench_c3_single.json
for w in c2 c1 c5 c5f32; do
  timeout 400 python bench.py --workload $w --steps 50 --warmup 5 --no-cpu --e2e-steps 2 > gpurun_out/bench_${w}.json 2> gpurun_out/bench_${w}.err || tail -3 gpurun_out/bench_${w}.err
  summ $w gpurun_out/bench_${w}.json
done
CMD="python bench.py --steps 6 --warmup 3 --no-cpu --e2e-steps 1 --variant tb2:128x16"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_tb2.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tb2 -s 4 -c 1 -o gpurun_out/prof_c3_tb2 -f $CMD > gpurun_out/ncu_full_tb2.log 2>&1
echo "ncu full tb2 exit $?"; tail -2 gpurun_out/ncu_full_tb2.log
CMD1="python bench.py --steps 6 --warmup 3 --no-cpu --e2e-steps 1 --tblock 1"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tma -s 4 -c 1 -o gpurun_out/prof_c3_tma -f $CMD1 > gpurun_out/ncu_full_tma.log 2>&1
echo "ncu full tma exit $?"; tail -2 gpurun_out/ncu_full_tma.log
