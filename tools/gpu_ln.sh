cd /root/repo
timeout 900 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "ln or modulate or layer" 2>&1 | tail -3
MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_vln2.so timeout 900 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "ln or modulate or layer" 2>&1 | tail -3
for i in 1 2; do
  echo "== previous (packed row, three unpack passes)"; MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_velprev.so timeout 300 python tools/bench_elem.py 2>&1 | tail -2 | head -1
  echo "== fp32-resident row, 3 CTAs/SM"; timeout 300 python tools/bench_elem.py 2>&1 | tail -2 | head -1
  echo "== fp32-resident row, 2 CTAs/SM"; MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_vln2.so timeout 300 python tools/bench_elem.py 2>&1 | tail -2 | head -1
done | tee gpurun_out/ln_resident_ab.txt
for tag in _velprev "" _vln2; do
  echo "== in-step libmdv2$tag.so"
  MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2$tag.so timeout 600 python bench.py --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > /tmp/b.json 2>/tmp/b.err
  python - <<'PY'
import json
d=json.loads(open('/tmp/b.json').read().strip().splitlines()[-1])
k=d['kernels']
print(d['ms_per_step'], d['clocks']['sm_mhz'], {n:(v['ms'],v['achieved']) for n,v in k.items() if v['share']>0.01})
PY
done | tee -a gpurun_out/ln_resident_ab.txt
