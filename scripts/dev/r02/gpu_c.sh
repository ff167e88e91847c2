#!/bin/bash
# r02 session C (1 GPU): tb2q with the TMA L2 prefetch
mkdir -p gpurun_out
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "reps", d["config"].get("rep_ms"), "parity", (d["config"].get("parity_check") or {}).get("bits_equal"), "clk", (d.get("clocks") or {}).get("sm_mhz"), (d.get("clocks") or {}).get("reasons"), (d.get("clocks") or {}).get("power_w_max"))
except Exception as e: print(sys.argv[1], "failed", e, open(sys.argv[2]).read()[-600:])
PY
}
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --e2e-steps 0 > gpurun_out/bench_c3_pf.json 2> gpurun_out/bench_c3_pf.err; echo "bench exit $?"; summ pf gpurun_out/bench_c3_pf.json
timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu --e2e-steps 0 --no-parity > gpurun_out/bench_c3_pf100.json 2>&1; summ pf-100 gpurun_out/bench_c3_pf100.json
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "temporal or repeat or full_size" 2>&1 | tail -3
