"""Timeline of the persistent attention kernel across item boundaries (TRACE build): the MMA issuer of query tile 0 stamps
every ring entry it consumes; prints the clocks between consecutive K/V tiles of CTA 5 — the steady period inside an
item and what an item boundary costs."""
import ctypes, os, sys
os.environ["MDV2_ATTN_PERSIST"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from magicdrive_v2_b200 import ops
STEPS, EV = 96, 12
lib = ops._lib
lib.mdv2_debug_attn_trace.argtypes = [ctypes.c_void_p]; lib.mdv2_debug_attn_trace.restype = None
n_seq, L, H, D = int(sys.argv[1]) if len(sys.argv) > 1 else 204, int(sys.argv[2]) if len(sys.argv) > 2 else 1600, 16, 72
C = H * D
qkv = torch.randn(n_seq * L, 3 * C, device="cuda", dtype=torch.bfloat16)
out = torch.empty(n_seq * L, C, device="cuda", dtype=torch.bfloat16)
buf = torch.zeros(20 * STEPS * EV + 8 + 256 * 8, dtype=torch.int64, device="cuda")
for it in range(2):
    if it == 1:
        lib.mdv2_debug_attn_trace(buf.data_ptr())
    ops.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], out, n_seq, L, L, H, D, D ** -0.5)
    torch.cuda.synchronize()
lib.mdv2_debug_attn_trace(None)
t = buf[:20 * STEPS * EV].view(20, STEPS, EV).cpu()
n_kt = (L + 127) // 128
for w in (17, 18):
    ev0 = t[w, :, 0]
    js = [j for j in range(STEPS) if ev0[j] > 0]
    print(f"mma warp {w}: ring entries stamped {js[:3]}..{js[-3:]} ({len(js)}); tiles per item {n_kt} (+1 Q entry)")
    prev = None
    row = []
    for j in js:
        e = t[w, j]
        d = int(e[0] - prev) if prev is not None else 0
        row.append(f"j{j}:{d}(P0w {int(e[1]-e[0])} I0 {int(e[2]-e[1])} P1w {int(e[3]-e[2])} I1 {int(e[4]-e[3])})")
        prev = int(e[0])
    for i in range(0, min(len(row), 45), 3):
        print("   " + "  ".join(row[i:i + 3]))

base = int(t[17, 1, 0])
print("softmax warps between passes (clk since the issuer's first tile): entry / pre-copy done / row sums exchanged / O complete / "
      "O read (o_empty) / stored / next pass ready;  then this warp's first-tile marks of the LAST pass for reference")
for w in (0, 4, 8, 12):
    for n in range(0, 4):
        e = t[w, 80 + n]
        print(f"  warp {w:2d} pass {n}: " + " ".join(f"{int(x) - base:8d}" for x in e[:7]))
for w in (17,):
    for j in (13, 15, 16, 27, 29):
        print(f"  mma{w} j={j}: " + " ".join(f"{int(x) - base:8d}" for x in t[w, j, :7]))
