#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/bench_attn.py --quick"
ITERS=1 timeout 200 $CMD > gpurun_out/plain_attn.log 2>&1 &&
ITERS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_sm100 -s 1 -c 1 \
    -o gpurun_out/prof_attn128_r01 -f $CMD > gpurun_out/ncu_attn.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_attn.log
