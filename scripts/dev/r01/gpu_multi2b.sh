#!/bin/bash
# 2-GPU pass: boundary launch on the high-priority stream (NEPTUNE_RT_SLAB_BOUNDARY_STREAM=1) - parity, then C3 at N=2 with and without
mkdir -p gpurun_out
T=tests/test_gpu_multi.py::test_slab_nccl_equals_single_domain
NEPTUNE_RT_SLAB_BOUNDARY_STREAM=1 timeout 300 python -m pytest -m gpu -x -q "$T[heat_star-shape4-9-2-0]" "$T[heat_star-shape8-6-2-0]" "$T[heat_box_2d9-shape9-8-2-0]" > gpurun_out/pytest_multi2b.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_multi2b.log
for sb in 1 0; do
f=gpurun_out/scale_c3_n2_sb$sb
NEPTUNE_RT_SLAB_BOUNDARY_STREAM=$sb timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --workload c3 --gpus 2 --steps 100 --warmup 6 --no-cpu --e2e-steps 0 > $f.json 2> $f.err
echo "c3 N=2 sb=$sb exit $?"; python - $f.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(d["n_gpus"], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "ms %.4f"%d["ms_per_step"], "launches", d["gpu_launches"])
except Exception as e: print("failed", e)
PY
done
