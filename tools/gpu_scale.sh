#!/bin/bash
# strong-scaling check on one 8-GPU box: N = 8, 4, 2
mkdir -p gpurun_out
nvidia-smi -L | wc -l
for n in ${NS:-8 4 2}; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "n=$n rc=$?"; tail -3 gpurun_out/bench_n$n.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n$n.json").read().strip().splitlines()[-1])
    print("N=$n", d["ms_per_step"], "sp_parity", d.get("sp_parity_rel_l2"), "e2e", d["e2e"]["value"], "launches", d["gpu_launches"], d["clocks"])
    tot=0
    for k,v in d["kernels"].items(): print("  ",k,v["calls"],v["ms"],v["achieved"],v["unit"]); tot+=v["ms"]
    print("   kernel sum", tot)
except Exception as e: print("parse failed", e)
PY
done
