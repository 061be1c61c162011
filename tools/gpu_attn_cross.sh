# persistent cross-attention kernel: parity, then A/B against the one-tile-per-CTA flavour on the same box
set -x
cd /root/repo
timeout 600 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "attention" 2>&1 | tail -5 | tee gpurun_out/cross_tests.txt
cat > /tmp/xb.py <<'PY'
import os, sys, torch
sys.path.insert(0, "/root/repo")
from tools.bench_attn import bench, accuracy
accuracy(L=333, Lk=62)
for n, lq in ((60, 5600), (204, 1600), (60, 1400), (60, 400)):
    bench(n, lq, 62, kind="cross", iters=10)
PY
for i in 1 2; do for c in 1 0; do echo "== MDV2_ATTN_CROSS=$c"; MDV2_ATTN_CROSS=$c timeout 300 python /tmp/xb.py 2>&1 | tail -5; done; done | tee gpurun_out/cross_ab.txt
