#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/tests_gpu.log


for n in 2; do
  if [ $n = 1 ]; then timeout 600 python bench.py --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  else timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; fi
  echo "n=$n rc=$?"; tail -3 gpurun_out/bench_n$n.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n$n.json").read().strip().splitlines()[-1])
    print("N=$n", d["ms_per_step"], "e2e", d["e2e"]["value"], "launches", d["gpu_launches"], d["clocks"])
    tot=0
    for k,v in d["kernels"].items(): print("  ",k,v["calls"],v["ms"],v["achieved"],v["unit"]); tot+=v["ms"]
    print("   kernel sum", tot)
except Exception as e: print("parse failed", e)
PY
done
