#!/bin/bash
# multi-GPU pass: parity across ranks, then the weak-scaling bench at N = 1, 2, (4, 8)
mkdir -p gpurun_out
nvidia-smi -L | tee gpurun_out/gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest multi exit $?"; tail -6 gpurun_out/pytest_multi.log
for n in "$@"; do
  if [ "$n" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 100 --warmup 5 --no-cpu > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 100 --warmup 5 --no-cpu > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
  echo "N=$n exit $?"; tail -1 gpurun_out/scale_n$n.json | cut -c1-400; tail -3 gpurun_out/scale_n$n.err
done
