#!/bin/bash
# r02 session K (1 GPU): tb2q with a pointwise input (C5): parity + variants on fp64 / fp32
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "pointwise or repeat or 1000 or jit or temporal" 2>&1 | tail -5
one() { timeout 300 python bench.py --workload $1 --variant $2 --steps 20 --warmup 5 --no-cpu --e2e-steps 0 $3 > gpurun_out/k_$1_${2/:/_}.json 2>gpurun_out/k_$1_${2/:/_}.err; python -c "
import json; d=json.loads(open('gpurun_out/k_$1_${2/:/_}.json').read().strip().splitlines()[-1]); print('$1 $2', '%.1f Gupd/s'%d['value'], d['config']['rep_ms'], d['config']['kernel_variant'], (d['config'].get('parity_check') or {}).get('bits_equal'), d['clocks']['reasons'])" || tail -3 gpurun_out/k_$1_${2/:/_}.err; }
one c5 tb2m:64x24 --no-parity
one c5 tb2q:64x24
one c5 tb2q:128x14 --no-parity
one c5f32 tb2m:128x14 --no-parity
one c5f32 tb2q:128x14
one c5f32 tb2q:64x24 --no-parity
