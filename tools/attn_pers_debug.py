"""Protocol debugging of the persistent attention kernel: runs one shape through the TRACE build, whose barrier waits
record a time-out (block, warp, source line, barrier index, parity, ring entry) instead of trapping."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from magicdrive_v2_b200 import ops

STEPS, EV = 96, 12
lib = ops._lib
lib.mdv2_debug_attn_trace.argtypes = [ctypes.c_void_p]
lib.mdv2_debug_attn_trace.restype = None
BARS = ["k_full0", "k_full1", "k_full2", "k_empty0", "k_empty1", "k_empty2", "v_full0", "v_full1", "v_full2",
        "v_empty0", "v_empty1", "v_empty2", "s_full00", "s_full01", "s_full10", "s_full11", "p_full00", "p_full01",
        "p_full10", "p_full11", "o_full0", "o_full1", "o_empty0", "o_empty1", "qt_full0", "qt_full1", "set_done"]


def run(n_seq, L, H=4, D=72, n_sets=1, gain=1.0):
    C = H * D
    g = torch.Generator(device="cuda").manual_seed(1)
    qkv = torch.randn(n_seq * L, 3 * C, device="cuda", dtype=torch.bfloat16, generator=g)
    qkv[:, :C] *= gain
    out = torch.zeros(n_seq * L, C, device="cuda", dtype=torch.bfloat16)
    kv_map = None
    if n_sets == 2:
        kv_map = torch.stack([(torch.arange(n_seq) + 1) % n_seq, (torch.arange(n_seq) + n_seq - 1) % n_seq], 1).int().contiguous().cuda()
    buf = torch.zeros(20 * STEPS * EV + 8 + 256 * 8, dtype=torch.int64, device="cuda")
    lib.mdv2_debug_attn_trace(buf.data_ptr())
    ops.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], out, n_seq, L, L, H, D, D ** -0.5, kv_map=kv_map)
    torch.cuda.synchronize()
    lib.mdv2_debug_attn_trace(None)
    t = buf[20 * STEPS * EV:].cpu()
    n = int(t[0])
    print(f"n_seq={n_seq} L={L} H={H} sets={n_sets} gain={gain}: {n} timed-out waits")
    only = int(os.environ.get("ONLY_BLOCK", "-1"))
    shown = 0
    for i in range(min(n, 256)):
        r = t[8 + i * 8: 16 + i * 8].tolist()
        if only >= 0 and r[0] != only:
            continue
        shown += 1
        if shown > 60:
            break
        print(f"  block {r[0]:3d} warp {r[1]:2d} line {r[2]:4d} bar {BARS[r[3]] if 0 <= r[3] < len(BARS) else r[3]} parity {r[4]} e {r[5]}")
    if n == 0:
        q, k, v = (qkv[:, i * C:(i + 1) * C].float().view(n_seq, L, H, D).transpose(1, 2) for i in range(3))
        ref = 0
        for s in range(n_sets):
            idx = kv_map[:, s].long() if kv_map is not None else torch.arange(n_seq, device="cuda")
            ref = ref + (torch.softmax(q @ k[idx].transpose(-1, -2) * D ** -0.5, dim=-1) @ v[idx]).to(torch.bfloat16).float()
        ref = ref.transpose(1, 2).reshape(n_seq * L, C)
        print("  rel L2 vs fp32 softmax:", ((out.float() - ref).norm() / ref.norm()).item())
    return n


if __name__ == "__main__":
    cases = ((6, 256), (6, 400), (2, 1408, 16), (12, 600, 16, 72, 2), (4, 1000, 4, 72, 1, 20.0))
    if len(sys.argv) > 1:
        cases = [cases[int(a)] for a in sys.argv[1:]]
    for args in cases:
        if run(*args):
            break
