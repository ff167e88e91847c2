#!/bin/bash
# iteration pass: parity, bench lines for the listed workloads, then launch list + full ncu capture of c3
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
for w in ${WORKLOADS:-c3 c2 c1 c5 c5f32}; do
  f=gpurun_out/bench_${w}.json
  timeout 400 python bench.py --workload $w --steps ${STEPS:-100} --warmup 5 --no-cpu --e2e-steps 3 $BENCH_ARGS > $f 2> gpurun_out/bench_${w}.err || tail -3 gpurun_out/bench_${w}.err
  python - "$w" $f <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%r["achieved"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "e2e %.2f"%d["e2e"]["value"], d["clocks"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
if [ -n "$NCU" ]; then
CMD="python bench.py --workload ${NCU_W:-c3} --steps 6 --warmup 3 --no-cpu --e2e-steps 1 $BENCH_ARGS"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_ -s 4 -c 1 -o gpurun_out/prof_${NCU_W:-c3} -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"; tail -2 gpurun_out/ncu_full.log
fi
