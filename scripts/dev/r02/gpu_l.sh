#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -q -x -k "reports or shape13 or shape19" > gpurun_out/pytest_multi_mode.log 2>&1; echo "exit $?"; tail -5 gpurun_out/pytest_multi_mode.log
