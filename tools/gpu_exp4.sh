#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_ops_gpu.py -q -x -k "gemm" 2>&1 | tail -3 | tee gpurun_out/ops_gemm.log
timeout 120 python tools/bench_gemm.py --model 2>&1 | tee gpurun_out/bench_gemm_model.log
timeout 120 python tools/bench_gemm.py --model 2>&1 | tee -a gpurun_out/bench_gemm_model.log
timeout 200 python tools/bench_gemm.py --sustained 2>&1 | tee gpurun_out/gemm_sustained.log
