#!/bin/bash
# 8-GPU pass: multi-GPU parity (world 4), then the weak-scaling bench at N = 8, 4, 2, 1 (C3) and N = 8 (C5 Jacobi)
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_multi.log
run() { # workload n tag
  local w=$1 n=$2 f=gpurun_out/scale_${1}_n$2
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --workload $w --gpus 1 --steps 100 --warmup 6 --no-cpu --e2e-steps 2 > $f.json 2> $f.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --workload $w --gpus $n --steps 100 --warmup 6 --no-cpu --e2e-steps 2 > $f.json 2> $f.err
  fi
  echo "$w N=$n exit $?"; python - $f.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(d["n_gpus"], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "ms %.4f"%d["ms_per_step"], "e2e %.2f"%d["e2e"]["value"], d["clocks"])
except Exception as e: print("failed", e)
PY
  tail -2 $f.err | cut -c1-300
}
for n in 8 4 2 1; do run c3 $n; done
run c5 8
run c5 1
