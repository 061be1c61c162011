#!/bin/bash
# N GPUs of one box (default 8): BASELINE.json configs[3] (65 frames, 424x800) through bench.py, and the configs[4]
# sampling loop (BrushNet / SDE-BrushNet denoiser, 424x800, CFG doubled) under sequence parallelism
N=${N:-8}; EXTRA=${EXTRA:-}
mkdir -p gpurun_out
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $RUN --master-port 29611 tools/bench_sampler.py --pad-sp $N --sampling-steps ${STEPS:-30} $EXTRA > gpurun_out/sampler_brushnet_n$N.json 2> gpurun_out/sampler_brushnet_n$N.err
echo "sampler rc=$?"; tail -1 gpurun_out/sampler_brushnet_n$N.json; tail -2 gpurun_out/sampler_brushnet_n$N.err
timeout 300 $RUN --master-port 29612 tools/bench_sampler.py --pad-sp $N --sde --sampling-steps 10 $EXTRA > gpurun_out/sampler_sde_n$N.json 2> gpurun_out/sampler_sde_n$N.err
echo "sde rc=$?"; tail -1 gpurun_out/sampler_sde_n$N.json; tail -2 gpurun_out/sampler_sde_n$N.err
if [ -z "$EXTRA" ]; then
timeout 400 $RUN --master-port 29613 bench.py --gpus $N --workload xl_65f_424x800 --steps 3 --warmup 3 --no-secondary > gpurun_out/bench_cfg4_n$N.json 2> gpurun_out/bench_cfg4_n$N.err
echo "cfg4 rc=$?"; tail -3 gpurun_out/bench_cfg4_n$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_cfg4_n$N.json").read().strip().splitlines()[-1])
print({k:d.get(k) for k in ("ms_per_step","value","value_conditions_cached","value_cuda_graph","sp_parity_rel_l2","cuda_graph_note")})
for k,v in d["kernels"].items(): print("  ",k,v["calls"],v["ms"],v["achieved"],v["unit"])
PY
fi
