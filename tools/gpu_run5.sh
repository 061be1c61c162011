#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/tests_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
echo "n1 rc=$?"; tail -5 gpurun_out/bench_n1.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_n1.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["e2e"]["value"], d["gpu_launches"], d["clocks"], d["step_frac_of_bf16_sustained"], d["step_frac_of_bf16_burst"])
for k,v in d["kernels"].items(): print("  ",k,v["calls"],v["ms"],v["achieved"],v["unit"])
PY
