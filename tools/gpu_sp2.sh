#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L | tee gpurun_out/gpus.txt
timeout 600 python -m pytest tests/test_sp_gpu.py -q -x 2>&1 | tail -8 | tee gpurun_out/sp_gpu_test.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
echo "rc=$?"; tail -c 2500 gpurun_out/bench_n2.json; tail -15 gpurun_out/bench_n2.err
