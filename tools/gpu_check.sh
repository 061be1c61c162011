#!/bin/bash
# round-2 check run: whole GPU test suite, the attention shapes (one-shot default, persistent opt-in), then the default bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/tests_gpu.log
MDV2_ATTN_PERSIST=1 timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_zz_no_alloc.py -k "attention or static" -x -q 2>&1 | tail -3 | tee gpurun_out/tests_attn_persist.log
for ps in 0 1; do echo "== MDV2_ATTN_PERSIST=$ps"; MDV2_ATTN_PERSIST=$ps timeout 300 python tools/bench_attn.py 2>&1 | grep "self\|xview\|cross"; done | tee gpurun_out/attn_bench.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02.json 2> gpurun_out/bench_r02.err
echo "bench rc=$?"; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r02.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','step_tflops_per_gpu','step_frac_of_bf16_sustained','gpu_launches','clocks')})
print('e2e',d['e2e']); print('roofline',{k:v for k,v in d['roofline'].items() if k!='traffic'})
for k,v in d['kernels'].items(): print(k, {x:v[x] for x in ('calls','ms','share','achieved','frac','unit')})
print('secondary', {k:d['secondary'][k] for k in ('value','ms_per_step','step_tflops')}); print(d.get('cpu_baseline')); print(d['secondary'].get('cpu_baseline'))
PY
tail -3 gpurun_out/bench_r02.err
