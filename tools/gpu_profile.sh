#!/bin/bash
# round-1 measurement run: full bench line, then ncu launch list and --set full captures of the top kernels
mkdir -p gpurun_out
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01.json 2> gpurun_out/bench_r01.err
echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_r01.json; tail -5 gpurun_out/bench_r01.err
CMD="python bench.py --steps 1 --warmup 3 --no-secondary --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s 6000 -c 3000 --csv \
    --log-file gpurun_out/launches_r01.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_sm100 -s 1500 -c 3 \
    -o gpurun_out/prof_gemm_r01 $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_sm100 -s 150 -c 3 \
    -o gpurun_out/prof_attn_r01 $CMD > gpurun_out/ncu_attn.log 2>&1
echo "attn rc=$?"
ls -la gpurun_out
