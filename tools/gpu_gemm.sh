#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_ops_gpu.py -q -x -k "gemm_tcgen05" 2>&1 | tail -15 | tee gpurun_out/ops_tc.log
timeout 300 python tools/bench_gemm.py 2>&1 | tail -20 | tee gpurun_out/bench_gemm.log
timeout 600 python -m pytest tests/test_model_gpu.py -q -x 2>&1 | tail -5 | tee gpurun_out/model.log
