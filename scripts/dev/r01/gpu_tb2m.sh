#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -12 gpurun_out/pytest_gpu.log
for v in "$@"; do
  for zc in ${ZCHUNKS:-0}; do
  f=gpurun_out/bench_c3_${v/:/_}_z${zc}.json
  timeout 300 python bench.py --steps ${STEPS:-100} --warmup 6 --no-cpu --e2e-steps 1 --variant $v --zchunk $zc > $f 2>> gpurun_out/variants.err
  python - "$v z$zc" $f <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%d["roofline"]["achieved"], "frac %.3f"%d["roofline"]["frac"], "ms/step %.4f"%d["ms_per_step"], "launches", d["gpu_launches"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
  done
done
tail -5 gpurun_out/variants.err
if [ -n "$NCU" ]; then
CMD="python bench.py --steps 8 --warmup 4 --no-cpu --e2e-steps 1 --variant $NCU"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tb2 -s 2 -c 1 -o gpurun_out/prof_c3_${NCU/:/_} -f $CMD > gpurun_out/ncu_full_tb2m.log 2>&1
echo "ncu full exit $?"; tail -2 gpurun_out/ncu_full_tb2m.log
fi
