#!/bin/bash
# r02 session F (N GPUs): p2p halo exchange — multi-GPU parity in both modes, slab timing by mode / ghost depth + trace
mkdir -p gpurun_out
N=${1:-2}
timeout 1200 python -m pytest tests/test_gpu_multi.py -q -x > gpurun_out/pytest_multi_p2p_n$N.log 2>&1; echo "pytest multi (p2p) exit $?"; tail -4 gpurun_out/pytest_multi_p2p_n$N.log
NEPTUNE_RT_HALO=nccl timeout 600 python -m pytest tests/test_gpu_multi.py -q -x -k "heat_star and not 1000" > gpurun_out/pytest_multi_nccl_n$N.log 2>&1; echo "pytest multi (nccl) exit $?"; tail -2 gpurun_out/pytest_multi_nccl_n$N.log
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["config"]["kernel_variant"], d["config"].get("halo_mode"), "%.1f Gupd/s"%d["value"], "ms/step %.4f"%d["ms_per_step"], "reps", d["config"].get("rep_ms"), "parity", (d["config"].get("parity_check") or {}).get("bits_equal"))
    for g in (d.get("slab_trace") or {}).get("groups", [])[:4]: print("    trace", g)
except Exception as e: print(sys.argv[1], "failed", e, open(sys.argv[2]).read()[-800:])
PY
}
run() { # name, extra args
  name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-cpu --e2e-steps 0 "$@" > gpurun_out/scale_${name}_n$N.json 2> gpurun_out/scale_${name}_n$N.err; echo "$name exit $?"; summ $name gpurun_out/scale_${name}_n$N.json
}
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --e2e-steps 0 --no-parity > gpurun_out/scale_c3_n1.json 2>/dev/null; summ n1 gpurun_out/scale_c3_n1.json
run c3_p2p_g2 --steps 20 --warmup 5 --ghost 2 --trace
run c3_p2p_g4 --steps 20 --warmup 5 --ghost 4 --trace
NEPTUNE_RT_HALO=nccl run c3_nccl_g2 --steps 20 --warmup 5 --ghost 2
run c3_p2p_g4_100 --steps 100 --warmup 5 --ghost 4 --no-parity
