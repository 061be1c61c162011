#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/tests_gpu.log
timeout 300 python tools/bench_attn.py 2>&1 | tail -20 | tee gpurun_out/bench_attn.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01.json 2> gpurun_out/bench_r01.err
echo "bench rc=$?"; tail -c 2500 gpurun_out/bench_r01.json; tail -5 gpurun_out/bench_r01.err
