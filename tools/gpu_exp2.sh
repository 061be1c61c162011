#!/bin/bash
mkdir -p gpurun_out
timeout 120 python tools/attn_trace.py 2>&1 | tee gpurun_out/attn_trace.log | tail -40
for bn in 0 128 64; do echo "MDV2_GEMM_BN=$bn"; MDV2_GEMM_BN=$bn timeout 120 python tools/bench_gemm.py --model; done 2>&1 | tee gpurun_out/bench_gemm_model.log
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:gemm_sm100 \
   -o gpurun_out/prof_gemm_pair_r01 -f python tools/bench_gemm.py --profile > gpurun_out/ncu_gemm_pair.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_gemm_pair.log
