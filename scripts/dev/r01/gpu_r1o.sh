#!/bin/bash
# session O (1 GPU): chunk count of the pipelined host path (e2e)
mkdir -p gpurun_out
for n in 16 32 48 64 128; do
  f=gpurun_out/bench_c3_e2e_chunks$n.json
  NEPTUNE_RT_HOST_CHUNKS=$n timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --e2e-steps 12 > $f 2>> gpurun_out/variants.err
  python - $n $f <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); e=d["e2e"]; print("chunks", sys.argv[1], "e2e %.3f Gupd/s"%e["value"], "ms %.2f"%e["ms_per_step"], "launches", d["e2e_gpu_launches"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
tail -3 gpurun_out/variants.err
