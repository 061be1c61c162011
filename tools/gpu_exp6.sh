#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_ops_gpu.py -q -x -k "attention_tcgen05 or scatter" 2>&1 | tail -2
ITERS=10 timeout 200 python tools/bench_attn.py 2>&1 | tail -6 | tee gpurun_out/bench_attn.log
