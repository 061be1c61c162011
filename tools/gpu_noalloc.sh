#!/bin/bash
# the no-allocation entry point: alone in a fresh process, then behind the whole GPU suite; then the default bench line
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/tests_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/tests_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err
echo "bench rc=$?"; tail -c 1500 gpurun_out/bench_r02.json; tail -3 gpurun_out/bench_r02.err
