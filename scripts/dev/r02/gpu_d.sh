#!/bin/bash
# r02 session D (1 GPU): tb2q with level-1 reads hoisted above the TMA wait
mkdir -p gpurun_out
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "reps", d["config"].get("rep_ms"), "parity", (d["config"].get("parity_check") or {}).get("bits_equal"), "clk", (d.get("clocks") or {}).get("sm_mhz"), (d.get("clocks") or {}).get("reasons"))
except Exception as e: print(sys.argv[1], "failed", e, open(sys.argv[2]).read()[-600:])
PY
}
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --e2e-steps 0 > gpurun_out/bench_c3_d.json 2> gpurun_out/bench_c3_d.err; echo "bench exit $?"; summ hoist gpurun_out/bench_c3_d.json
timeout 600 python bench.py --workload c5 --steps 20 --warmup 5 --no-cpu --e2e-steps 0 > gpurun_out/bench_c5_d.json 2> gpurun_out/bench_c5_d.err; summ c5 gpurun_out/bench_c5_d.json
timeout 600 python bench.py --workload c5f32 --steps 20 --warmup 5 --no-cpu --e2e-steps 0 > gpurun_out/bench_c5f32_d.json 2> gpurun_out/bench_c5f32_d.err; summ c5f32 gpurun_out/bench_c5f32_d.json
timeout 600 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu --e2e-steps 0 > gpurun_out/bench_c4_d.json 2> gpurun_out/bench_c4_d.err; summ c4 gpurun_out/bench_c4_d.json
timeout 600 python bench.py --workload c2 --steps 20 --warmup 5 --no-cpu --e2e-steps 0 > gpurun_out/bench_c2_d.json 2> gpurun_out/bench_c2_d.err; summ c2 gpurun_out/bench_c2_d.json
timeout 600 python bench.py --workload c1 --steps 100 --warmup 5 --no-cpu --e2e-steps 0 > gpurun_out/bench_c1_d.json 2> gpurun_out/bench_c1_d.err; summ c1 gpurun_out/bench_c1_d.json
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4
