#!/bin/bash
mkdir -p gpurun_out
for pf in 0 1 2 3; do
  echo "== MDV2_ATTN_PF=$pf"
  MDV2_ATTN_PF=$pf timeout 300 python -m pytest tests/test_ops_gpu.py -q -x -k "attention_tcgen05" 2>&1 | tail -1
  MDV2_ATTN_PF=$pf ITERS=10 timeout 200 python tools/bench_attn.py --quick 2>&1 | tail -2
done 2>&1 | tee gpurun_out/attn_pf.log
timeout 120 python tools/attn_trace.py 2>&1 | tail -22 | head -10 | tee gpurun_out/attn_trace.log
