# cross-attention kernel: parity, then timing experiments (stale K/V, no stores, stale Q) to find what bounds an item
cd /root/repo
timeout 600 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "attention" 2>&1 | tail -3 | tee gpurun_out/cross_tests.txt
cat > /tmp/xb.py <<'PY'
import os, sys, torch
sys.path.insert(0, "/root/repo")
from tools.bench_attn import bench
for n, lq in ((60, 5600), (204, 1600)):
    bench(n, lq, 62, kind="cross", iters=10)
PY
for d in 0 1 2 3 4 5 7; do echo "== MDV2_ATTN_DBG=$d"; MDV2_ATTN_DBG=$d timeout 300 python /tmp/xb.py 2>&1 | tail -2; done | tee gpurun_out/cross_exp.txt
