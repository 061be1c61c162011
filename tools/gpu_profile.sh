#!/bin/bash
# measurement run: GPU tests, full bench line, then ncu launch list (one bracketed step) and --set full captures
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/tests_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r01.json 2> gpurun_out/bench_r01.err
echo "bench rc=$?"; tail -c 2500 gpurun_out/bench_r01.json; tail -5 gpurun_out/bench_r01.err
CMD="python bench.py --steps 1 --warmup 3 --no-secondary --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches_r01.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    --kernel-name-base demangled -k "regex:attn_sm100_kernel<80, 128>" -s 2 -c 2 \
    -o gpurun_out/prof_attn_r01 -f $CMD > gpurun_out/ncu_attn.log 2>&1
echo "attn rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:temporal_attn -s 2 -c 1 -o gpurun_out/prof_temporal_r01 -f $CMD > gpurun_out/ncu_temporal.log 2>&1
echo "temporal rc=$?"
ls -la gpurun_out
