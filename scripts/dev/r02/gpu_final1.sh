#!/bin/bash
# r02 final 1-GPU session: full GPU suite, smoke, default bench line (+ reference arm), launch list, ncu of the default kernels
mkdir -p gpurun_out
export NEPTUNE_B200_JIT_CACHE=/tmp/jitcache
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu_final.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/smoke_final.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke_final.log
timeout 900 python bench.py > gpurun_out/bench_default_final.json 2> gpurun_out/bench_default_final.err; echo "bench exit $?"; tail -2 gpurun_out/bench_default_final.err
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_default_20.json 2> gpurun_out/bench_default_20.err; echo "bench20 exit $?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_final.json 2> gpurun_out/bench_reference_final.err; echo "ref exit $?"
python - <<'PY'
import json
for f in ["bench_default_final","bench_default_20"]:
    try:
        d=json.loads(open("gpurun_out/%s.json"%f).read().strip().splitlines()[-1]); r=d["roofline"]
        print(f, d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "frac %.3f"%r["frac"], "parity", d["config"]["parity_check"]["bits_equal"], "e2e", (d.get("e2e") or {}).get("value"), "cpu", (d.get("cpu_baseline") or {}).get("value"), "single", (r.get("single_step_kernel") or {}).get("value"), (r.get("single_step_kernel") or {}).get("frac"), d["clocks"])
    except Exception as e: print(f, "failed", e)
PY
tail -c 500 gpurun_out/bench_reference_final.json
CMD="python bench.py --steps 8 --warmup 3 --reps 1 --no-cpu --e2e-steps 1 --no-parity"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_default_final.csv $CMD > gpurun_out/ncu_list_final.log 2>&1; echo "ncu list exit $?"
CMD2="python bench.py --steps 4 --warmup 3 --reps 1 --no-cpu --e2e-steps 0 --no-parity"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tb2q -s 2 -c 1 -o gpurun_out/prof_c3_tb2q_final $CMD2 > gpurun_out/ncu_tb2q_final.log 2>&1; echo "ncu tb2q exit $?"
timeout 900 ncu --set full --clock-control none -k regex:sweep_tma -s 1 -c 1 -o gpurun_out/prof_c3_tma_final python bench.py --tblock 1 --steps 4 --warmup 3 --reps 1 --no-cpu --e2e-steps 0 --no-parity > gpurun_out/ncu_tma_final.log 2>&1; echo "ncu tma exit $?"
