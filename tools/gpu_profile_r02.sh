#!/bin/bash
# round-2 ncu evidence: launch list of ONE bracketed step of the bench command, --set full captures of the dominant
# kernels (the command runs plain first; ncu only if it exits 0)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-secondary --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/bench_plain_for_ncu.json 2> gpurun_out/bench_plain_for_ncu.err &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file gpurun_out/launches_r02.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:attn_fast_kernel -c 2 -o gpurun_out/prof_attn_r02 -f $CMD > gpurun_out/ncu_attn.log 2>&1
echo "attn rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k regex:gemm_sm100_kernel -s 200 -c 8 -o gpurun_out/prof_gemm_r02 -f $CMD > gpurun_out/ncu_gemm.log 2>&1
echo "gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off \
    -k "regex:attn_cross_kernel|temporal_attn_smem_kernel|ln_modulate_kernel" -s 30 -c 6 -o gpurun_out/prof_hbm_r02 -f $CMD > gpurun_out/ncu_hbm.log 2>&1
echo "hbm kernels rc=$?"
ls -la gpurun_out | tail -8
