#!/bin/bash
# session T (1 GPU, end of round 1): full GPU suite, smoke, complete default bench line + reference arm, launch list
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default_full.json 2> gpurun_out/bench_default_full.err; echo "bench exit $?"; tail -3 gpurun_out/bench_default_full.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%r["achieved"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "e2e", (d.get("e2e") or {}).get("value"), "cpu", (d.get("cpu_baseline") or {}).get("value"), "single", (r.get("single_step_kernel") or {}).get("value"))
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
summ default gpurun_out/bench_default_full.json
tail -1 gpurun_out/bench_reference.json | cut -c1-400
CMD="python bench.py --steps 8 --warmup 3 --no-cpu --e2e-steps 1"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_default.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "ncu list exit $?"
