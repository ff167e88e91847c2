#!/bin/bash
# r02 session H8 (8 GPUs): BASELINE configs as worded — C3 weak scaling at N = 8, C4 1024^3 over 8 GPUs, C5 768^3 over 8 GPUs
# (fp64 + fp32, 1000 iterations, temporal depth 1 / 2 / 4), e2e + host-link probe at N = 8
mkdir -p gpurun_out
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    e=d.get("e2e") or {}
    print(sys.argv[1], d["config"]["kernel_variant"], d["config"].get("halo_mode"), "%.1f Gupd/s"%d["value"], "ms/step %.5f"%d["ms_per_step"], "spread %.3f"%(d["config"].get("rep_spread") or 0), "parity", (d["config"].get("parity_check") or {}).get("bits_equal"), "e2e", e.get("value"), (e.get("host_link_probe") or {}).get("GBps_each_way_per_rank"), "clk", (d.get("clocks") or {}).get("sm_mhz"), (d.get("clocks") or {}).get("reasons"))
    for g in (d.get("slab_trace") or {}).get("groups", [])[:4]: print("    trace", g)
except Exception as e: print(sys.argv[1], "failed", e, open(sys.argv[2]).read()[-800:])
PY
}
run() { N=$1; name=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-cpu "$@" > gpurun_out/scale_${name}_n$N.json 2> gpurun_out/scale_${name}_n$N.err; echo "$name n$N exit $?"; summ "$name-n$N" gpurun_out/scale_${name}_n$N.json; }
run 8 c3 --steps 20 --warmup 5 --e2e-steps 3 --link-probe --trace
run 8 c3_100 --steps 100 --warmup 5 --e2e-steps 0 --no-parity
NEPTUNE_RT_HALO=nccl run 8 c3_nccl_g2 --steps 20 --warmup 5 --e2e-steps 0 --ghost 2 --no-parity
run 8 c4 --workload c4 --steps 20 --warmup 3 --e2e-steps 0
for d in 1 2 4; do
  run 8 c5g_f64_d$d --workload c5g --steps 1000 --warmup 10 --reps 3 --e2e-steps 0 --tblock $d
  run 8 c5g_f32_d$d --workload c5gf32 --steps 1000 --warmup 10 --reps 3 --e2e-steps 0 --tblock $d --no-parity
done
run 8 c5g_f64_d8 --workload c5g --steps 1000 --warmup 10 --reps 3 --e2e-steps 0 --tblock 8 --no-parity
