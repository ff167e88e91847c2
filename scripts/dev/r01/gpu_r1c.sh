#!/bin/bash
# 2-GPU pass: full parity (single + multi), weak-scaling bench at N=1,2 with the temporally blocked
# slab path, then launch list + full ncu capture of the temporally blocked kernel on one GPU
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log
summ() { python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=d["roofline"]
    print(sys.argv[1], d["config"]["kernel_variant"], "%.1f Gupd/s"%d["value"], "%.0f GB/s"%r["achieved"], "frac %.3f"%r["frac"], "ms %.4f"%d["ms_per_step"], "e2e %.2f"%d["e2e"]["value"], "single", (r.get("single_step_kernel") or {}).get("value"), "launches", d["gpu_launches"], d["clocks"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
}
timeout 600 python bench.py --gpus 1 --steps 100 --warmup 6 --no-cpu --e2e-steps 2 > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err; summ N=1 gpurun_out/scale_n1.json; tail -3 gpurun_out/scale_n1.err
for n in 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 100 --warmup 6 --no-cpu --e2e-steps 2 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  summ N=$n gpurun_out/scale_n$n.json; tail -3 gpurun_out/scale_n$n.err
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 100 --warmup 6 --no-cpu --e2e-steps 2 --tblock 1 > gpurun_out/scale_n${n}_tb1.json 2> gpurun_out/scale_n${n}_tb1.err
  summ N=$n-tb1 gpurun_out/scale_n${n}_tb1.json; tail -3 gpurun_out/scale_n${n}_tb1.err
done
CMD="python bench.py --steps 8 --warmup 4 --no-cpu --e2e-steps 1"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/launches_tb2.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sweep_tb2 -s 2 -c 1 -o gpurun_out/prof_c3_tb2 -f $CMD > gpurun_out/ncu_full_tb2.log 2>&1
echo "ncu full tb2 exit $?"; tail -2 gpurun_out/ncu_full_tb2.log
