for cfg in "MDV2_ATTN_DBG=0" "MDV2_ATTN_DBG=32"; do echo "== $cfg"; env $cfg timeout 200 python tools/bench_attn.py 2>&1 | grep "self\|xview"; env $cfg timeout 400 python bench.py --steps 3 --warmup 3 --no-secondary --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('step ms', d['ms_per_step'], d['clocks']['sm_mhz']); [print('  ',k,v['ms'],v['achieved']) for k,v in d['kernels'].items() if v['ms']>50]"; done
