#!/bin/bash
# session L (1 GPU): full GPU suite with the golden-fixture and full-size property tests
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -22 gpurun_out/pytest_gpu.log
