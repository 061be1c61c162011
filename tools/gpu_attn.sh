#!/bin/bash
# attention iteration run: parity tests of the attention kernels, per-shape numbers, CTA timeline, library yard-sticks
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv | tee gpurun_out/attn_gpu.txt
timeout 600 python -m pytest tests/test_ops_gpu.py -k "attention" -x -q 2>&1 | tail -5 | tee gpurun_out/attn_tests.log
for pf in 0 2 3; do
  echo "== MDV2_ATTN_PF=$pf" | tee -a gpurun_out/attn_bench.log
  MDV2_ATTN_PF=$pf timeout 300 python tools/bench_attn.py $([ $pf != 2 ] && echo --quick) 2>&1 | tee -a gpurun_out/attn_bench.log
done
timeout 300 python tools/attn_trace.py 2>&1 | tee gpurun_out/attn_trace.log
timeout 600 python tools/bench_attn.py --yardstick 2>&1 | tee gpurun_out/yardstick_attn.log
