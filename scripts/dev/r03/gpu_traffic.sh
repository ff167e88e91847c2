#!/bin/bash
# r03 (1 GPU): bench line of every BASELINE config, then one `ncu --set full` capture of its dominant kernel (DRAM traffic
# per launch for profiles/traffic.json, counters for profiles/r03_<w>_ncu.txt), and the launch list of the default command.
mkdir -p gpurun_out
export NEPTUNE_B200_JIT_CACHE=/tmp/jitcache
for w in c1 c2 c4 c5 c5f32; do
  timeout 900 python bench.py --workload $w --cpu-seconds 6 > gpurun_out/r03_${w}_bench.json 2> gpurun_out/r03_${w}_bench.err; echo "bench $w exit $?"
  python - gpurun_out/r03_${w}_bench.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "sustained %.1f"%d["sustained"]["value"], "frac %.3f"%r["frac"], "parity", d["config"]["parity_check"]["bits_equal"], "e2e %.2f"%d["e2e"]["value"])
except Exception as e: print("failed", e, open(sys.argv[1]).read()[-400:])
PY
  src=""; [ $w = c5f32 ] && src="--import-source on"
  timeout 900 ncu --set full --clock-control none $src -k regex:sweep_ -s 2 -c 1 -o gpurun_out/r03_prof_${w} \
    python bench.py --workload $w --steps 4 --warmup 3 --reps 1 --no-cpu --e2e-steps 0 --no-parity --sustained-seconds 0 > gpurun_out/r03_ncu_${w}.log 2>&1; echo "ncu $w exit $?"
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r03_c3_default_launches.csv \
  python bench.py --steps 8 --warmup 3 --reps 1 --no-cpu --e2e-steps 1 --no-parity --sustained-seconds 0 > gpurun_out/r03_ncu_list.log 2>&1; echo "ncu list exit $?"
ls -la gpurun_out | head -40
