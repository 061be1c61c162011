# attention kernels: shared-space (LDS / STS) vs generic accesses to the exchange / Q-copy / schedule records, same box
cd /root/repo
timeout 900 python -m pytest tests/test_ops_gpu.py -x -q -m gpu -k "attention" 2>&1 | tail -3 | tee gpurun_out/attn_lds_tests.txt
for i in 1 2; do
  echo "== one-shot, generic (previous)"; MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_vfg.so timeout 300 python tools/bench_attn.py 2>&1 | grep "self\|xview"
  echo "== one-shot, shared-space"; timeout 300 python tools/bench_attn.py 2>&1 | grep "self\|xview"
done | tee gpurun_out/attn_lds_ab.txt
for i in 1 2; do
  echo "== persistent, generic (previous)"; MDV2_ATTN_PERSIST=1 MDV2_LIB_FILE=/root/repo/magicdrive_v2_b200/libmdv2_vpg.so timeout 300 python tools/bench_attn.py 2>&1 | grep "self\|xview"
  echo "== persistent, shared-space"; MDV2_ATTN_PERSIST=1 timeout 300 python tools/bench_attn.py 2>&1 | grep "self\|xview"
done | tee -a gpurun_out/attn_lds_ab.txt
