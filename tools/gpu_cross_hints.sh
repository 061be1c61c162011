# cross-attention: ring / store-path variants (same box)
cd /root/repo
cat > /tmp/xb.py <<'PY'
import os, sys, torch
sys.path.insert(0, "/root/repo")
from tools.bench_attn import bench
for n, lq in ((60, 5600), (204, 1600), (60, 400)):
    bench(n, lq, 62, kind="cross", iters=10)
PY
for tag in "" _vr62 _vlsu _vlsu62; do
  echo "== libmdv2$tag.so"
  MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2$tag.so timeout 300 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "attention_tcgen05_cross" 2>&1 | tail -1
  for i in 1 2; do MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2$tag.so timeout 300 python /tmp/xb.py 2>&1 | tail -3; done
done | tee gpurun_out/cross_variants.txt
