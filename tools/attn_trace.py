"""Timeline of one attention CTA (debug build hook mdv2_debug_attn_trace): clock64 stamps per warp / K-V step / event.
softmax warps (fast kernel, per K/V tile): 0 S^hf ready, 1 chunk 0 done, 2 chunk 1 done, 3 P stored + arrive
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
for it in range(2):
    if it == 1:
        lib.mdv2_debug_attn_trace(buf.data_ptr())
    ops.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], out, n_seq, L, L, H, D, D ** -0.5)
    torch.cuda.synchronize()
lib.mdv2_debug_attn_trace(None)
t = buf.cpu()
base = int(t[t > 0].min())
import statistics
# softmax warps: index = half-step (j*2 + half); phase durations averaged over half-steps 16..80
# fast kernel (attn_fast_sm100.cu): warps 0..15 are softmax warps (t, hf) = (w>>3, (w>>2)&1); index = K/V tile
# events: 0 S^hf ready, 1 chunk 0 done (ld + exp + P store issued), 2 chunk 1 done, 3 P stored + arrive
for w in range(0, 16):
    d = {k: [] for k in ("chunk0", "chunk1", "st+arrive", "wait_S", "period")}
    for g in range(8, 40):
        e = t[w, g]
        d["chunk0"].append(int(e[1] - e[0])); d["chunk1"].append(int(e[2] - e[1])); d["st+arrive"].append(int(e[3] - e[2]))
        d["wait_S"].append(int(t[w, g + 1, 0] - e[3])); d["period"].append(int(t[w, g + 1, 0] - e[0]))
    print(f"warp {w:2d} (tile {w >> 3}, half {(w >> 2) & 1}, quarter {w & 3}; per K/V tile): " +
          "  ".join(f"{k} {statistics.mean(v):7.1f}" for k, v in d.items()))
# MMA warps (17: tile 0, 18: tile 1): per K/V tile j: 0 top (K/V ready), 1+2h P^h ready, 2+2h PV^h + QK^h(next) issued
for mw in (17, 18):
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
    for mw in (17, 18):
        print(f"j={j} mma{mw}:", " ".join(f"{int(x) - base:7d}" for x in t[mw, j, :5]))
    print(f"j={j} mma17 loop top / V_j seen / K_j+1 seen:", " ".join(f"{int(x) - base:7d}" for x in (t[17, j, 5], t[17, j, 6], t[17, j, 0])))
    print(f"j={j} tma: K_j stage free / V_j stage free (loads issued):", " ".join(f"{int(x) - base:7d}" for x in t[16, j, :2]),
          f"  K_j+2, V_j+2:", " ".join(f"{int(x) - base:7d}" for x in t[16, j + 2, :2]))
    for w in (0, 4, 8, 12):       # quarter 0: (tile 0, half 0), (0, 1), (1, 0), (1, 1)
        print(f"j={j} warp {w:2d}:", " ".join(f"{int(x) - base:7d}" for x in t[w, j, :4]))
