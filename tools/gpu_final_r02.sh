#!/bin/bash
# end-of-round check on one B200: GPU tests, smoke, library yard-sticks, the default bench line and the reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee gpurun_out/tests_gpu.log
timeout 600 python __graft_entry__.py smoke 2>&1 | tail -3 | tee gpurun_out/smoke.log
timeout 600 python tools/bench_attn.py --yardstick 2>&1 | tee gpurun_out/yardstick_attn.log
timeout 400 python tools/bench_gemm.py --sustained 2>&1 | tail -6 | tee gpurun_out/yardstick_gemm.log
timeout 900 python bench.py > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err
echo "bench rc=$?"; tail -2 gpurun_out/bench_r02.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_r02.json 2> gpurun_out/bench_ref_r02.err
echo "ref rc=$?"; tail -c 700 gpurun_out/bench_ref_r02.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r02.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','ms_per_step','step_tflops_per_gpu','step_frac_of_bf16_sustained','gpu_launches','clocks','value_conditions_cached','value_cuda_graph')})
print('e2e',d['e2e']); print('roofline',{k:v for k,v in d['roofline'].items() if k!='traffic'})
for k,v in d['kernels'].items(): print(k, {x:v[x] for x in ('calls','ms','share','achieved','frac','unit')})
s=d['secondary']; print('secondary', {k:s.get(k) for k in ('value','ms_per_step','step_tflops','value_cuda_graph')}); print(d.get('cpu_baseline')); print(s.get('cpu_baseline'))
PY
