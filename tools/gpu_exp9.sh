#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_rflow_gpu.py -q -x 2>&1 | tail -3
timeout 300 python -m pytest tests/test_ops_gpu.py -q -x -k "attention" 2>&1 | tail -3
echo "--- ping-pong on"; ITERS=10 timeout 200 python tools/bench_attn.py 2>&1 | tail -6 | tee gpurun_out/bench_attn.log
echo "--- ping-pong off"; MDV2_ATTN_DBG=2 ITERS=10 timeout 200 python tools/bench_attn.py 2>&1 | tail -5 | head -3
timeout 120 python tools/attn_trace.py 2>&1 | tail -22 | tee gpurun_out/attn_trace.log
CMD="python bench.py --steps 1 --warmup 3 --no-secondary --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:gemm_sm100_kernel -s 60 -c 6 -o gpurun_out/prof_gemm_r01 -f $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:attn_sm100_kernel -s 1 -c 1 -o gpurun_out/prof_cross_r01 -f $CMD > gpurun_out/ncu_cross.log 2>&1
echo "cross rc=$?"
