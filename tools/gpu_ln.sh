cd /root/repo
timeout 900 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "ln or modulate or layer" 2>&1 | tail -3
for i in 1 2; do
  echo "== previous"; MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_velprev.so timeout 300 python tools/bench_elem.py 2>&1 | tail -2
  echo "== packed fp32 pairs"; timeout 300 python tools/bench_elem.py 2>&1 | tail -2
done | tee gpurun_out/ln_packed_ab.txt
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_model_gpu.py -x -q -m gpu 2>&1 | tail -3
