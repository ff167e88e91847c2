#!/bin/bash
# 4-GPU pass: merged boundary launch (middle ranks have a neighbour on both sides) - parity subset, then C3 at N=4
mkdir -p gpurun_out
T=tests/test_gpu_multi.py::test_slab_nccl_equals_single_domain
timeout 600 python -m pytest -m gpu -x -q "$T[heat_box_2d9-shape1-7-None-0]" "$T[heat_star-shape4-9-2-0]" "$T[heat_star-shape8-6-2-0]" "$T[heat_box_2d9-shape9-8-2-0]" "$T[jacobi_star-shape10-8-2-0]" > gpurun_out/pytest_multi4b.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_multi4b.log
f=gpurun_out/scale_c3_n4
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511 bench.py --workload c3 --gpus 4 --steps 100 --warmup 6 --no-cpu --e2e-steps 2 > $f.json 2> $f.err
echo "c3 N=4 exit $?"; python - $f.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(d["n_gpus"], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "ms %.4f"%d["ms_per_step"], "launches", d["gpu_launches"], d["clocks"])
except Exception as e: print("failed", e)
PY
tail -2 $f.err | cut -c1-300
