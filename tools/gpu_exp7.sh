#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_ops_gpu.py -q -x -k "attention_tcgen05" 2>&1 | tail -1
ITERS=10 timeout 200 python tools/bench_attn.py 2>&1 | tail -6 | tee gpurun_out/bench_attn.log
timeout 120 python tools/attn_trace.py 2>&1 | tail -22 | head -10 | tee gpurun_out/attn_trace.log
CMD="python bench.py --steps 1 --warmup 3 --no-secondary --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k 'regex:attn_sm100_kernel.*\)128' -s 2 -c 1 -o gpurun_out/prof_attn_r01 -f $CMD > gpurun_out/ncu_attn.log 2>&1
echo "attn rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k 'regex:gemm_sm100_kernel.*\)192' -s 3 -c 2 -o gpurun_out/prof_gemm_r01 -f $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "gemm rc=$?"
ls -la gpurun_out/*.ncu-rep
