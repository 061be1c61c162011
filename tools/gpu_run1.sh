#!/bin/bash
# first GPU bring-up: safe kernels first, then the tcgen05 GEMM in its own process (bounded)
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv | tee gpurun_out/gpu.txt
timeout 600 python -m pytest tests/test_ops_gpu.py -q -k "not tcgen05" 2>&1 | tail -25 | tee gpurun_out/ops_simt.log
timeout 300 python -m pytest tests/test_ops_gpu.py -q -x -k "tcgen05" 2>&1 | tail -40 | tee gpurun_out/ops_tc.log
timeout 600 python -m pytest tests/test_model_gpu.py -q -s -k "fp32" 2>&1 | tail -25 | tee gpurun_out/model_fp32.log
timeout 600 python -m pytest tests/test_model_gpu.py -q -s -k "not fp32" 2>&1 | tail -25 | tee gpurun_out/model_bf16.log
timeout 300 python tools/bench_gemm.py 2>&1 | tail -20 | tee gpurun_out/bench_gemm.log
