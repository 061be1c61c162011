#!/bin/bash
# N-GPU check of the fused exchanges: SP tests (3 exchange modes), then the bench (MDV2_SP_FUSED=0 run: UNFUSED=1)
N=${N:-2}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_sp_gpu.py -q -x 2>&1 | tail -3
run() {  # $1 = tag, env in front
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_n${N}_$1.json 2> gpurun_out/bench_n${N}_$1.err
  echo "$1 rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_n${N}_$1.json").read().strip().splitlines()[-1])
    print("$1", d["n_gpus"], "gpus", round(d["ms_per_step"], 1), "ms/step", {k: round(v["ms"], 1) for k, v in d["kernels"].items()})
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/bench_n${N}_$1.err").read()[-1500:])
PY
}
run fused
if [ -n "$UNFUSED" ]; then MDV2_SP_FUSED=0 run unfused; fi
