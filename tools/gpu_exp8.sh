#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_rflow_gpu.py -q -x 2>&1 | tail -5
CMD="python bench.py --steps 1 --warmup 3 --no-secondary --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:attn_sm100_kernel -c 1 -o gpurun_out/prof_attn_r01 -f $CMD > gpurun_out/ncu_attn.log 2>&1
echo "attn rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:gemm_sm100_kernel -s 4 -c 4 -o gpurun_out/prof_gemm_r01 -f $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:temporal_attn_smem_kernel -c 1 -o gpurun_out/prof_temporal_r01 -f $CMD > gpurun_out/ncu_temporal.log 2>&1
echo "temporal rc=$?"
ls -la gpurun_out/*.ncu-rep
