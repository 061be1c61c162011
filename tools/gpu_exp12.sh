#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_ops_gpu.py -q -x -k "attention" 2>&1 | tail -2
echo "--- RS=1 pingpong"; ITERS=10 timeout 200 python tools/bench_attn.py 2>&1 | tail -6 | tee gpurun_out/bench_attn.log
echo "--- RS=1 no pingpong"; MDV2_ATTN_DBG=2 ITERS=10 timeout 200 python tools/bench_attn.py 2>&1 | tail -5 | head -2
echo "--- RS=1 PF=0 pingpong"; MDV2_ATTN_PF=0 ITERS=10 timeout 200 python tools/bench_attn.py 2>&1 | tail -5 | head -2
timeout 120 python tools/attn_trace.py 2>&1 | tail -22 | head -10 | tee gpurun_out/attn_trace.log
export MDV2_ATTN_RS=2
timeout 300 python -m pytest tests/test_ops_gpu.py -q -x -k "attention" 2>&1 | tail -2
echo "--- RS=2 PF=2 pingpong"; ITERS=10 timeout 200 python tools/bench_attn.py 2>&1 | tail -5 | head -2
echo "--- RS=2 PF=0 pingpong"; MDV2_ATTN_PF=0 ITERS=10 timeout 200 python tools/bench_attn.py 2>&1 | tail -5 | head -2
echo "--- RS=2 PF=2 no pingpong"; MDV2_ATTN_DBG=2 ITERS=10 timeout 200 python tools/bench_attn.py 2>&1 | tail -5 | head -2
timeout 120 python tools/attn_trace.py 2>&1 | tail -22 | head -10 | tee gpurun_out/attn_trace_rs2.log
