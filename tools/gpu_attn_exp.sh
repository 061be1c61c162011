for cfg in "MDV2_ATTN_DBG=0" "MDV2_ATTN_PERSIST=0"; do echo "== $cfg"; env $cfg timeout 200 python tools/bench_attn.py 2>&1 | grep "self\|xview"; done
MDV2_ATTN_PERSIST=1 timeout 200 python tools/attn_trace.py 2>&1 | grep -v "^warp  [1-9]\|^warp 1[0-5]" | head -12
