# GEMM: elected-lane MMA issue + lean qkv epilogue (new lib) vs the previous build (libmdv2_vold.so); BN = 216 vs 144
cd /root/repo
timeout 900 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "gemm" 2>&1 | tail -3 | tee gpurun_out/gemm216_tests.txt
for i in 1 2; do
  echo "== old lib"; MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_vold.so timeout 300 python tools/bench_gemm.py --model 2>&1 | tail -5
  for v in 1 0; do echo "== new lib MDV2_GEMM_QKV216=$v"; MDV2_GEMM_QKV216=$v timeout 300 python tools/bench_gemm.py --model 2>&1 | tail -5; done
done | tee gpurun_out/gemm216_ab.txt
echo "== old lib"; MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_vold.so timeout 300 python tools/bench_gemm.py --sustained 2>&1 | tail -4 | tee gpurun_out/gemm_sustained_old.txt
echo "== new lib"; timeout 300 python tools/bench_gemm.py --sustained 2>&1 | tail -4 | tee gpurun_out/gemm_sustained_new.txt
