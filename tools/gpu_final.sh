#!/bin/bash
# late-round check on one B200: GPU parity tests, the default bench line, the BrushNet sampling loop (configs[4] shape)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/tests_gpu.log
echo "tests rc=${PIPESTATUS[0]}"; tail -5 gpurun_out/tests_gpu.log
timeout 420 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench rc=$?"; head -c 900 gpurun_out/bench.json; echo; tail -3 gpurun_out/bench.err
timeout 240 python tools/bench_sampler.py > gpurun_out/bench_sampler.json 2> gpurun_out/bench_sampler.err
echo "sampler rc=$?"; cat gpurun_out/bench_sampler.json; tail -3 gpurun_out/bench_sampler.err
timeout 180 python tools/bench_sampler.py --sde --sampling-steps 10 > gpurun_out/bench_sampler_sde.json 2> gpurun_out/bench_sampler_sde.err
echo "sde rc=$?"; cat gpurun_out/bench_sampler_sde.json; tail -3 gpurun_out/bench_sampler_sde.err
