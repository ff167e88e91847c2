#!/bin/bash
# N-GPU pass: multi-GPU parity (middle ranks have two neighbours), then the weak-scaling bench
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_multi.log
for n in "$@"; do
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 100 --warmup 6 --no-cpu --e2e-steps 3 > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 100 --warmup 6 --no-cpu --e2e-steps 3 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
  echo "N=$n exit $?"; python - gpurun_out/scale_n$n.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(d["n_gpus"], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "ms %.4f"%d["ms_per_step"], "e2e %.2f"%d["e2e"]["value"], d["clocks"])
except Exception as e: print("failed", e)
PY
  tail -2 gpurun_out/scale_n$n.err | cut -c1-300
done
