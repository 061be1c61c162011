#!/bin/bash
# checkpoint run on 2 GPUs: all GPU tests (incl. NCCL SP), bench at N=1 and N=2
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/tests_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
echo "n1 rc=$?"; tail -c 1500 gpurun_out/bench_n1.json; tail -5 gpurun_out/bench_n1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
echo "n2 rc=$?"; tail -c 2500 gpurun_out/bench_n2.json; tail -5 gpurun_out/bench_n2.err
