#!/bin/bash
# First GPU pass: parity tests, smoke, bench (all variants), ncu launch list + full capture.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 100 --warmup 5 --omp-context > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err; echo "bench exit $?"; cat gpurun_out/bench_c3.json; tail -3 gpurun_out/bench_c3.err
for v in direct tma:64x16 tma:128x8 tma:128x16 tma:64x32; do
  timeout 300 python bench.py --steps 50 --warmup 5 --no-cpu --e2e-steps 1 --variant $v > gpurun_out/bench_c3_${v/:/_}.json 2>> gpurun_out/variants.err
  python - "$v" gpurun_out/bench_c3_${v/:/_}.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); print(sys.argv[1], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%d["roofline"]["achieved"], "frac %.3f"%d["roofline"]["frac"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
CMD="python bench.py --steps 5 --warmup 3 --no-cpu --e2e-steps 1"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_ -s 4 -c 2 -o gpurun_out/prof_c3 -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"; tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out
