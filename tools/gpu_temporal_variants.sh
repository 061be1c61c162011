# short-clip temporal kernel: geometry variants (columns per CTA / threads / CTAs per SM), same box
cd /root/repo
for i in 1 2; do for tag in "" _vg1 _vg1b _vg3 _vg2t; do
  echo "== libmdv2$tag.so"
  MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2$tag.so timeout 300 python tools/bench_elem.py 2>&1 | tail -1
done; done | tee gpurun_out/temporal_variants.txt
for tag in _vg1 _vg3; do
  MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2$tag.so timeout 600 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k temporal 2>&1 | tail -2
done | tee gpurun_out/temporal_variants_tests.txt
bash tools/gpu_cross_ncu.sh
