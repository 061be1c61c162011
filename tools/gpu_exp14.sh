#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/tests_gpu.log
for P in 0 72 66; do
  MDV2_GEMM_PAIRS=$P timeout 120 python tools/gemm_pairs_probe.py
  MDV2_GEMM_PAIRS=$P timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_op_read_hit_rate.pct \
      --clock-control none -k regex:gemm_sm100_kernel -s 5 -c 1 python tools/gemm_pairs_probe.py 2>&1 | grep -E "dram__|gpu__time|hit_rate"
done
