#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py -m gpu -x -q -k "gemm" 2>&1 | tail -4
for t in 1 0; do echo "== MDV2_GEMM_TAIL=$t"; MDV2_GEMM_TAIL=$t timeout 300 python tools/bench_gemm.py --sustained 2>&1 | tail -12; done | tee gpurun_out/gemm_tail.log
