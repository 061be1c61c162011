cd /root/repo
timeout 900 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "gemm or attention" 2>&1 | tail -3 | tee gpurun_out/mix_tests.txt
for i in 1 2; do
  echo "== previous build"; MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_vprev.so timeout 300 python tools/bench_gemm.py --model 2>&1 | tail -5
  echo "== shared-space staging accesses"; timeout 300 python tools/bench_gemm.py --model 2>&1 | tail -5
done | tee gpurun_out/gemm_lds_ab.txt
for tag in _vprev ""; do
  echo "== in-step libmdv2$tag.so"
  MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2$tag.so timeout 600 python bench.py --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > /tmp/b.json 2>/tmp/b.err
  python - <<'PY'
import json
d=json.loads(open('/tmp/b.json').read().strip().splitlines()[-1])
k=d['kernels']
print(d['ms_per_step'], d['clocks']['sm_mhz'], {n:(v['ms'],v['achieved']) for n,v in k.items() if v['share']>0.01})
PY
done | tee -a gpurun_out/gemm_lds_ab.txt
