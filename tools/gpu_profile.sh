#!/bin/bash
# measurement run: GPU tests, full default bench line, reference arm, ncu launch list (one bracketed step) and
# --set full captures of the dominant kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/tests_gpu.log
timeout 900 python bench.py > gpurun_out/bench_r01.json 2> gpurun_out/bench_r01.err
echo "bench rc=$?"; tail -c 1800 gpurun_out/bench_r01.json; tail -3 gpurun_out/bench_r01.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_r01.json 2> gpurun_out/bench_ref_r01.err
echo "ref rc=$?"; tail -c 600 gpurun_out/bench_ref_r01.json
CMD="python bench.py --steps 1 --warmup 3 --no-secondary --no-cpu-baseline"
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches_r01.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
# (-k matches the function base name; the first attention launch of the bracketed step is the 128-key self-attention,
#  the second the 64-key cross-attention; the first ~170 GEMM launches are the small condition-side projections)
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:attn_sm100_kernel -c 2 -o gpurun_out/prof_attn_r01 -f $CMD > gpurun_out/ncu_attn.log 2>&1
echo "attn rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:gemm_sm100_kernel -s 200 -c 8 -o gpurun_out/prof_gemm_r01 -f $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:"temporal_attn|ln_modulate" -s 2 -c 3 -o gpurun_out/prof_hbm_r01 -f $CMD > gpurun_out/ncu_hbm.log 2>&1
echo "hbm rc=$?"
ls -la gpurun_out | tail -12
