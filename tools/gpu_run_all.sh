#!/bin/bash
# bring-up run: each group in its own bounded process
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv | tee gpurun_out/gpu.txt
nproc | tee -a gpurun_out/gpu.txt
timeout 600 python -m pytest tests/test_ops_gpu.py -q -k "not tcgen05" 2>&1 | tail -25 | tee gpurun_out/ops_simt.log
timeout 300 python -m pytest tests/test_ops_gpu.py -q -k "gemm_tcgen05" 2>&1 | tail -40 | tee gpurun_out/ops_tc.log
timeout 300 python -m pytest tests/test_ops_gpu.py -q -k "attention_tcgen05" 2>&1 | tail -40 | tee gpurun_out/ops_attn_tc.log
timeout 600 python -m pytest tests/test_model_gpu.py -q -s 2>&1 | tail -40 | tee gpurun_out/model.log
timeout 300 python tools/bench_gemm.py 2>&1 | tail -20 | tee gpurun_out/bench_gemm.log
