#!/bin/bash
# r02 session A (1 GPU): full GPU suite, smoke, tb2p vs tb2q on C3, ncu of tb2q
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "reps", d["config"].get("rep_ms"), "parity", (d["config"].get("parity_check") or {}).get("bits_equal"), "e2e", (d.get("e2e") or {}).get("value"), "single", (r.get("single_step_kernel") or {}).get("value"))
except Exception as e: print(sys.argv[1], "failed", e, open(sys.argv[2]).read()[-400:])
PY
}
for v in tb2p:128x14 tb2q:128x14; do
  timeout 600 python bench.py --variant $v --steps 20 --warmup 5 --no-cpu --e2e-steps 0 > gpurun_out/bench_c3_${v/:/_}.json 2> gpurun_out/bench_c3_${v/:/_}.err; echo "bench $v exit $?"
  summ $v gpurun_out/bench_c3_${v/:/_}.json
done
timeout 600 python bench.py --variant tb2q:128x14 --steps 100 --warmup 5 --no-cpu --e2e-steps 0 --no-parity > gpurun_out/bench_c3_tb2q_100.json 2>&1; summ tb2q-100 gpurun_out/bench_c3_tb2q_100.json
CMD="python bench.py --variant tb2q:128x14 --steps 4 --warmup 3 --reps 1 --no-cpu --e2e-steps 0 --no-parity"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tb2q -s 2 -c 1 -o gpurun_out/prof_c3_tb2q $CMD > gpurun_out/ncu_tb2q.log 2>&1; echo "ncu exit $?"; tail -2 gpurun_out/ncu_tb2q.log
ls -la gpurun_out | head -30
