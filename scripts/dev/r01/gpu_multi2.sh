#!/bin/bash
# 2-GPU pass: multi-GPU parity incl. rank-2 slabs with two ghost rows (sweep_rows2), weak-scaling lines for C2 / C3
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi2.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_multi2.log
for w in c2 c3; do
  f=gpurun_out/scale_${w}_n2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --workload $w --gpus 2 --steps 60 --warmup 6 --no-cpu --e2e-steps 0 > $f.json 2> $f.err
  echo "$w N=2 exit $?"; python - $f.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(d["n_gpus"], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "ms %.4f"%d["ms_per_step"], d["config"]["decomposition"])
except Exception as e: print("failed", e)
PY
  tail -2 $f.err | cut -c1-300
done
