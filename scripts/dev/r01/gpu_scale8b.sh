#!/bin/bash
# 8-GPU box, bench only: C3 weak scaling at N = 8 and 2 with the tb2p default (N = 1 comes from the 1-GPU run)
mkdir -p gpurun_out
run() { # workload n
  local w=$1 n=$2 f=gpurun_out/scale_${1}_n$2
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --workload $w --gpus $n --steps 100 --warmup 6 --no-cpu --e2e-steps 2 > $f.json 2> $f.err
  echo "$w N=$n exit $?"; python - $f.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(d["n_gpus"], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "ms %.4f"%d["ms_per_step"], "e2e %.2f"%d["e2e"]["value"], d["clocks"])
except Exception as e: print("failed", e)
PY
  tail -2 $f.err | cut -c1-300
}
run c3 8
run c3 2
