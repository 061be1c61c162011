timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_zz_no_alloc.py -k "attention or static" -x -q 2>&1 | tail -3
for cfg in "MDV2_ATTN_DBG=0" "MDV2_ATTN_DBG=16"; do echo "== $cfg"; env $cfg timeout 200 python tools/bench_attn.py 2>&1 | grep "self\|xview"; done
timeout 200 python tools/attn_cta_trace.py 2>&1 | grep "L= 1600\|L= 5600\|L=  512" | cut -c1-330
