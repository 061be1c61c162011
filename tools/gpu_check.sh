#!/bin/bash
# round-2 check run: whole GPU test suite, then the default bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/tests_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err
echo "bench rc=$?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r02.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','step_tflops_per_gpu','step_frac_of_bf16_sustained','gpu_launches','clocks')})
print('e2e',d['e2e']); print('roofline',d['roofline'])
for k,v in d['kernels'].items(): print(k, {x:v[x] for x in ('calls','ms','share','achieved','frac','unit')})
print('secondary', {k:d['secondary'][k] for k in ('value','ms_per_step','step_tflops')}); print(d.get('cpu_baseline')); print(d['secondary'].get('cpu_baseline'))
PY
tail -3 gpurun_out/bench_r02.err
