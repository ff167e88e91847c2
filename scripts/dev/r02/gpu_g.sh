#!/bin/bash
# r02 session G (2 GPUs): why is a tb2q launch 5 % slower inside a 2-rank job?  (hypothesis: peer-access-enabled allocations)
mkdir -p gpurun_out
N=2
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["config"]["kernel_variant"], d["config"].get("halo_mode"), "%.1f Gupd/s"%d["value"], "ms/step %.4f"%d["ms_per_step"], "reps", d["config"].get("rep_ms"))
    for g in (d.get("slab_trace") or {}).get("groups", [])[:4]: print("    trace", g, "dur %.4f" % (g[4]-g[1]))
except Exception as e: print(sys.argv[1], "failed", e, open(sys.argv[2]).read()[-800:])
PY
}
run() { name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-cpu --e2e-steps 0 --no-parity "$@" > gpurun_out/g_${name}.json 2> gpurun_out/g_${name}.err; echo "$name exit $?"; summ $name gpurun_out/g_${name}.json; }
NEPTUNE_RT_HALO=nccl NCCL_P2P_DISABLE=1 run nccl_nop2p --steps 20 --warmup 5 --ghost 4 --trace
NEPTUNE_RT_HALO=nccl run nccl --steps 20 --warmup 5 --ghost 4 --trace
run p2p --steps 20 --warmup 5 --ghost 4 --trace
# two independent single-GPU runs at once (no NCCL, no peer access): does a second busy GPU matter?
(CUDA_VISIBLE_DEVICES=1 timeout 300 python bench.py --steps 100 --warmup 5 --reps 3 --no-cpu --e2e-steps 0 --no-parity > gpurun_out/g_solo1.json 2>/dev/null &)
CUDA_VISIBLE_DEVICES=0 timeout 300 python bench.py --steps 100 --warmup 5 --reps 3 --no-cpu --e2e-steps 0 --no-parity > gpurun_out/g_solo0.json 2>/dev/null; sleep 5
summ solo0 gpurun_out/g_solo0.json; summ solo1 gpurun_out/g_solo1.json
