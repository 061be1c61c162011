#!/bin/bash
# persistent attention kernel: parity tests, then per-shape numbers against the one-shot kernel and the libraries
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_zz_no_alloc.py -k "attention or static" -x -q 2>&1 | tail -15 | tee gpurun_out/attn_tests.log
for persist in 1 0; do
  echo "== MDV2_ATTN_PERSIST=$persist" | tee -a gpurun_out/attn_bench.log
  MDV2_ATTN_PERSIST=$persist timeout 300 python tools/bench_attn.py 2>&1 | tee -a gpurun_out/attn_bench.log
done
