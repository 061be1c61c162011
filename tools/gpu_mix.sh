cd /root/repo
timeout 900 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "gemm or attention" 2>&1 | tail -3 | tee gpurun_out/mix_tests.txt
cat > /tmp/xb.py <<'PY'
import os, sys, torch
sys.path.insert(0, "/root/repo")
from tools.bench_attn import bench
for n, lq in ((60, 5600), (204, 1600), (60, 1400), (60, 400)):
    bench(n, lq, 62, kind="cross", iters=10)
PY
for i in 1 2; do for t in 1 0; do echo "== MDV2_ATTN_CROSS_TEAMS=$t"; MDV2_ATTN_CROSS_TEAMS=$t timeout 300 python /tmp/xb.py 2>&1 | tail -4; done; done | tee gpurun_out/cross_teams.txt
for i in 1 2; do
  echo "== previous gemm"; MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_vgprev.so timeout 300 python tools/bench_gemm.py --model 2>&1 | tail -5
  echo "== residual prefetched under the main loop"; timeout 300 python tools/bench_gemm.py --model 2>&1 | tail -5
done | tee gpurun_out/gemm_resprefetch_ab.txt
