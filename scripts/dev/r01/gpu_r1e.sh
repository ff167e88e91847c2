#!/bin/bash
# round-1 session E: reduce tests, complete default bench line (cpu_baseline in), reference arm, launch list
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/gpu.txt; nproc >> gpurun_out/gpu.txt
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -6 gpurun_out/pytest_gpu.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default_full.json 2> gpurun_out/bench_default_full.err; echo "bench exit $?"; cat gpurun_out/bench_default_full.json; tail -3 gpurun_out/bench_default_full.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"; cat gpurun_out/bench_reference.json; tail -3 gpurun_out/bench_reference.err
CMD="python bench.py --steps 8 --warmup 3 --no-cpu --e2e-steps 1"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_default.csv $CMD > gpurun_out/ncu_list.log 2>&1; echo "ncu list exit $?"
