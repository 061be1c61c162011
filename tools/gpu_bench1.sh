#!/bin/bash
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench1.err
echo "rc=$?"; tail -c 6000 gpurun_out/bench1.json; tail -20 gpurun_out/bench1.err
