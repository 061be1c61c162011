# in-step A/B of the attention kernel's polynomial-exponential fraction (power-capped step, same box)
cd /root/repo
for pf in 3 0 2 4 3; do
  echo "== MDV2_ATTN_PF=$pf"
  MDV2_ATTN_PF=$pf timeout 600 python bench.py --steps 3 --warmup 3 --no-secondary --no-cpu-baseline > /tmp/b.json 2>/tmp/b.err
  python - <<'PY'
import json
d=json.loads(open('/tmp/b.json').read().strip().splitlines()[-1])
k=d['kernels']
print(d['ms_per_step'], d['clocks']['sm_mhz'], {n:(v['ms'],v['achieved']) for n,v in k.items() if v['share']>0.01})
PY
done | tee gpurun_out/pf_instep.txt
