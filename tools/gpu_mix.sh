cd /root/repo
timeout 900 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "temporal" 2>&1 | tail -3 | tee gpurun_out/mix_tests.txt
for i in 1 2 3; do
  echo "== previous build"; MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_vlprev.so timeout 300 python tools/bench_elem.py 2>&1 | tail -1
  echo "== packed pairs in the score and P.V phases"; timeout 300 python tools/bench_elem.py 2>&1 | tail -1
done | tee gpurun_out/temporal_ffma2_ab.txt
