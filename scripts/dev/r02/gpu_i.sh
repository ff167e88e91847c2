#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --workload c4 --steps 5 --warmup 3 --reps 2 --no-cpu --e2e-steps 0 > gpurun_out/i_c4.json 2> gpurun_out/i_c4.err; echo "c4 exit $?"; tail -c 600 gpurun_out/i_c4.json; tail -3 gpurun_out/i_c4.err
timeout 600 python bench.py --workload c5g --steps 20 --warmup 3 --reps 2 --no-cpu --e2e-steps 0 > gpurun_out/i_c5g.json 2> gpurun_out/i_c5g.err; echo "c5g exit $?"; python -c "
import json; d=json.loads(open('gpurun_out/i_c5g.json').read().strip().splitlines()[-1]); print(d['value'], d['config']['parity_check']['bits_equal'], d['config']['kernel_variant'])"
