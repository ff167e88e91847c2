#!/bin/bash
# session Q (1 GPU): tb2p as the default C3 kernel: full suite, default bench, ncu full capture, launch list
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default_full.json 2> gpurun_out/bench_default_full.err; echo "bench exit $?"; tail -2 gpurun_out/bench_default_full.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_default_full.json").read().strip().splitlines()[-1]); r=d["roofline"]
print(d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "frac %.3f"%r["frac"], "e2e", d["e2e"]["value"], "cpu", d["cpu_baseline"]["value"], d["cpu_baseline"].get("context_openmp"), "single", r["single_step_kernel"]["value"])
PY
CMD="python bench.py --steps 8 --warmup 4 --no-cpu --e2e-steps 1"
timeout 300 $CMD > gpurun_out/plain_tb2p.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tb2p -s 2 -c 1 -o gpurun_out/prof_c3_tb2p -f $CMD > gpurun_out/ncu_full_tb2p.log 2>&1
echo "ncu full exit $?"; tail -2 gpurun_out/ncu_full_tb2p.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_default.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "ncu list exit $?"
