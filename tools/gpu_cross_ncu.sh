#!/bin/bash
# one --set full capture of the persistent cross-attention kernel (60 x 5600 queries onto 62 tokens)
mkdir -p gpurun_out
cat > /tmp/xb.py <<'PY'
import sys
sys.path.insert(0, "/root/repo")
from tools.bench_attn import bench
bench(60, 5600, 62, kind="cross", iters=3)
PY
python /tmp/xb.py > gpurun_out/cross_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:attn_cross_kernel -s 2 -c 1 \
    -o gpurun_out/prof_attn_cross -f python /tmp/xb.py > gpurun_out/ncu_cross.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_cross.log; cat gpurun_out/cross_plain.log
