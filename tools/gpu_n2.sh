#!/bin/bash
# two GPUs of one box: the sequence-parallel parity tests (skipped on a 1-GPU box), then the bench line at N = 2
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_sp_gpu.py -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/tests_sp_gpu_n2.log
NS=2 bash tools/gpu_scale.sh 2>&1 | tee gpurun_out/scale_n2.log
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_n2.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("value_conditions_cached","value_cuda_graph","cuda_graph_note","sp_parity_rel_l2")}); print(d["config"].get("sp_exchange"))
PY
