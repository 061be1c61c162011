#!/bin/bash
# --set full captures of the five in-model GEMM flavours (stand-alone, M = 84 000)
mkdir -p gpurun_out
python tools/bench_gemm.py --model > gpurun_out/gemm_model_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:gemm_sm100_kernel -o gpurun_out/prof_gemm_model -f python tools/bench_gemm.py --profile > gpurun_out/ncu_gemm_model.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_gemm_model.log; tail -5 gpurun_out/gemm_model_plain.log
