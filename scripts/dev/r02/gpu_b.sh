#!/bin/bash
# r02 session B (1 GPU): full GPU suite (JIT, reference fixtures, memref), smoke, default bench with padded tb2q, ncu
mkdir -p gpurun_out
export NEPTUNE_B200_JIT_CACHE=/tmp/jitcache
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "reps", d["config"].get("rep_ms"), "parity", (d["config"].get("parity_check") or {}).get("bits_equal"), "e2e", (d.get("e2e") or {}).get("value"), "single", (r.get("single_step_kernel") or {}).get("value"))
except Exception as e: print(sys.argv[1], "failed", e, open(sys.argv[2]).read()[-600:])
PY
}
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --e2e-steps 0 > gpurun_out/bench_c3_auto.json 2> gpurun_out/bench_c3_auto.err; echo "bench exit $?"; summ auto gpurun_out/bench_c3_auto.json
timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu --e2e-steps 0 --no-parity > gpurun_out/bench_c3_auto100.json 2>&1; summ auto-100 gpurun_out/bench_c3_auto100.json
CMD="python bench.py --steps 4 --warmup 3 --reps 1 --no-cpu --e2e-steps 0 --no-parity"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tb2q -s 2 -c 1 -o gpurun_out/prof_c3_tb2q_v2 $CMD > gpurun_out/ncu_tb2q_v2.log 2>&1; echo "ncu exit $?"; tail -2 gpurun_out/ncu_tb2q_v2.log
