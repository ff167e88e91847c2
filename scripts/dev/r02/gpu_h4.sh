#!/bin/bash
# r02 session H4 (4 GPUs): multi-GPU parity with middle ranks (two neighbours) in p2p mode, C3 at N = 4 and N = 2
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multi.py -q -x > gpurun_out/pytest_multi_p2p_n4.log 2>&1; echo "pytest multi n4 exit $?"; tail -4 gpurun_out/pytest_multi_p2p_n4.log
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
    print(sys.argv[1], d["config"]["kernel_variant"], d["config"].get("halo_mode"), "%.1f Gupd/s"%d["value"], "ms/step %.4f"%d["ms_per_step"], "reps", d["config"].get("rep_ms"), "parity", (d["config"].get("parity_check") or {}).get("bits_equal"), "e2e", (d.get("e2e") or {}).get("value"))
    for g in (d.get("slab_trace") or {}).get("groups", [])[:4]: print("    trace", g)
except Exception as e: print(sys.argv[1], "failed", e, open(sys.argv[2]).read()[-800:])
PY
}
run() { N=$1; name=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-cpu "$@" > gpurun_out/scale_${name}_n$N.json 2> gpurun_out/scale_${name}_n$N.err; echo "$name n$N exit $?"; summ "$name-n$N" gpurun_out/scale_${name}_n$N.json; }
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu --e2e-steps 0 --no-parity > gpurun_out/scale_c3_n1.json 2>/dev/null; summ n1 gpurun_out/scale_c3_n1.json
run 4 c3 --steps 20 --warmup 5 --e2e-steps 0 --trace
run 2 c3 --steps 20 --warmup 5 --e2e-steps 0
NEPTUNE_RT_HALO=nccl run 4 c3_nccl_g2 --steps 20 --warmup 5 --e2e-steps 0 --ghost 2
