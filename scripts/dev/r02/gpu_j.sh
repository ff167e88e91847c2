#!/bin/bash
# r02 session J (1 GPU): sweep_tma ring read with LDS vs generic LD.E, A/B in one session
mkdir -p gpurun_out
one() { timeout 300 python bench.py --workload $1 --tblock 1 --variant $2 --steps $3 --warmup 5 --no-cpu --e2e-steps 0 --no-parity > gpurun_out/j_$1_$4.json 2>gpurun_out/j_$1_$4.err; python -c "
import json; d=json.loads(open('gpurun_out/j_$1_$4.json').read().strip().splitlines()[-1]); print('$1 $2', '%.1f Gupd/s'%d['value'], d['config']['rep_ms'], d['config']['kernel_variant'], d['clocks']['reasons'])"; }
one c3 tma:64x16 20 lds
one c3 tma:64x16g 20 gen
one c3 tma:64x16 20 lds2
one c3 tma:64x16g 20 gen2
one c5 tma:64x16 10 lds
one c5 tma:64x16g 10 gen
one c4 tma:64x8 10 lds
one c4 tma:64x8g 10 gen
