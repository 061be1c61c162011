#!/bin/bash
mkdir -p gpurun_out
for dbg in 0 2; do
  echo "== MDV2_ATTN_DBG=$dbg"
  MDV2_ATTN_DBG=$dbg MDV2_ATTN_PF=2 ITERS=10 timeout 200 python tools/bench_attn.py --quick 2>&1 | tail -1
  MDV2_ATTN_DBG=$dbg MDV2_ATTN_PF=2 timeout 120 python tools/attn_trace.py 2>&1 | tail -28 | grep -v "^j=1[34]" | cut -c1-200
done 2>&1 | tee gpurun_out/attn_dbg.log
