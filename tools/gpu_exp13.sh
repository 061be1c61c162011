#!/bin/bash
export MDV2_ATTN_RS=2
timeout 120 python tools/attn_trace.py 2>&1 | tail -38 | tee gpurun_out/attn_trace_rs2.log
echo ---- no pingpong
MDV2_ATTN_DBG=2 timeout 120 python tools/attn_trace.py 2>&1 | tail -38 | head -18
