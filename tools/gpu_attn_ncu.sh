#!/bin/bash
# one --set full capture of the spatial attention kernel (S = 5600 self-attention, one launch)
mkdir -p gpurun_out
export ITERS=1
python tools/bench_attn.py --quick > gpurun_out/attn_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_fast_kernel -s 2 -c 1 \
    -o gpurun_out/prof_attn_fast -f python tools/bench_attn.py --quick > gpurun_out/ncu_attn.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_attn.log; cat gpurun_out/attn_plain.log | grep self
