#!/bin/bash
# session M (1 GPU): grouped z-march (sweep_tb2g)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "temporal_blocking_bit_exact" > gpurun_out/pytest_tb2g.log 2>&1; echo "pytest exit $?"; tail -5 gpurun_out/pytest_tb2g.log
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%r["achieved"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "launches", d["gpu_launches"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
: > gpurun_out/variants.err
for v in ${VARIANTS:-tb2g:128x14 tb2g:64x24 tb2m:128x14}; do
  for zc in ${ZCHUNKS:-0}; do
  f=gpurun_out/bench_c3_${v/:/_}_z$zc.json
  timeout 300 python bench.py --steps 100 --warmup 6 --no-cpu --e2e-steps 0 --variant $v --zchunk $zc > $f 2>> gpurun_out/variants.err; summ "c3 $v z$zc" $f
  done
done
tail -5 gpurun_out/variants.err
