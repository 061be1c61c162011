cd /root/repo
timeout 900 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "gemm" 2>&1 | tail -3 | tee gpurun_out/mix_tests.txt
for i in 1 2 3; do
  echo "== previous build"; MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_vgprev.so timeout 300 python tools/bench_gemm.py --model 2>&1 | tail -5
  echo "== staged vectors and bias elements read ahead"; timeout 300 python tools/bench_gemm.py --model 2>&1 | tail -5
done | tee gpurun_out/gemm_hoist_ab.txt
