"""Timeline of one attention CTA (debug build hook mdv2_debug_attn_trace): clock64 stamps per warp / K-V step / event.
softmax warps: 0 loop top, 1 S ready, 2 S in registers, 3 max+rescale done, 4 exp+pack done, 5 P stored + arrive
MMA warp:      0 K/V ready, 1 P0 ready, 2 PV0+QK0' issued, 3 P1 ready, 4 PV1+QK1' issued"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from magicdrive_v2_b200 import ops

STEPS, EV = 96, 12
lib = ops._lib
lib.mdv2_debug_attn_trace.argtypes = [ctypes.c_void_p]
lib.mdv2_debug_attn_trace.restype = None
n_seq, L, H, D = 60, 5600, 16, 72
C = H * D
qkv = torch.randn(n_seq * L, 3 * C, device="cuda", dtype=torch.bfloat16)
out = torch.empty(n_seq * L, C, device="cuda", dtype=torch.bfloat16)
buf = torch.zeros(20, STEPS, EV, dtype=torch.int64, device="cuda")
RS = 1
for it in range(2):
    if it == 1:
        lib.mdv2_debug_attn_trace(buf.data_ptr())
    ops.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], out, n_seq, L, L, H, D, D ** -0.5)
    torch.cuda.synchronize()
lib.mdv2_debug_attn_trace(None)
t = buf.cpu()
base = int(t[t > 0].min())
import statistics
tw = int(os.environ.get("MDV2_ATTN_DBG", "0")) & 1      # row of the TMA warp holds the per-MMA issue stamps of group (0,0)
mw = 1 - tw
# softmax warps: index = half-step (j*2 + half); phase durations averaged over half-steps 16..80
for w in ((2, 3, 4, 5, 6, 7, 8, 9) if RS == 1 else (2, 3, 6, 7, 10, 11, 14, 15)):
    d = {k: [] for k in ("wait_S", "ld", "max", "exp", "st", "period")}
    for g in range(16, 80):
        e = t[w, g]
        d["wait_S"].append(int(e[1] - e[0])); d["ld"].append(int(e[2] - e[1])); d["max"].append(int(e[3] - e[2]))
        d["exp"].append(int(e[4] - e[3])); d["st"].append(int(e[5] - e[4])); d["period"].append(int(t[w, g + 1, 0] - e[0]))
    print(f"warp {w} (per half-step): " + "  ".join(f"{k} {statistics.mean(v):7.1f}" for k, v in d.items()))
    if RS == 2:   # exp = token wait + exponentials; st = exchange + P store
        tok = [int(t[w, g, 6] - t[w, g, 3]) for g in range(16, 80)]
        ex = [int(t[w, g, 4] - t[w, g, 6]) for g in range(16, 80)]
        xc = [int(t[w, g, 7] - t[w, g, 4]) for g in range(16, 80)]
        sp = [int(t[w, g, 5] - t[w, g, 7]) for g in range(16, 80)]
        print(f"        token wait {statistics.mean(tok):7.1f}  exp {statistics.mean(ex):7.1f}  exchange {statistics.mean(xc):7.1f}"
              f"  P store {statistics.mean(sp):7.1f}")
# MMA warps (1: tile 0, 10: tile 1): per K/V tile j: 0 top (K/V ready), 1+2h P^h ready, 2+2h PV^h + QK^h(next) issued
for mw in ((1, 10) if RS == 1 else (1, 18)):
    names = ["P0", "I0", "P1", "I1"]
    dd = {n: [] for n in names}
    per, top = [], []
    for j in range(8, 40):
        e = t[mw, j]
        for i in range(1, 5):
            dd[names[i - 1]].append(int(e[i] - e[i - 1]))
        per.append(int(t[mw, j + 1, 0] - e[0]))
        top.append(int(t[mw, j + 1, 0] - e[4]))
    print(f"mma warp {mw} (per K/V tile): " + "  ".join(f"{k} {statistics.mean(v):6.1f}" for k, v in dd.items()) +
          f"  top {statistics.mean(top):6.1f}  period {statistics.mean(per):7.1f}")
j0 = int(os.environ.get("J0", "12"))
for j in range(j0, j0 + 2):
    for mw in ((1, 10) if RS == 1 else (1, 18)):
        print(f"j={j} mma{mw}:", " ".join(f"{int(x) - base:7d}" for x in t[mw, j, :5]))
    for hf in (0, 1):
        for w, name in ((2, "sm0"), (6, "sm1")):
            print(f"j={j} h={hf} {name}:", " ".join(f"{int(x) - base:7d}" for x in t[w, j * 2 + hf, :6]))
