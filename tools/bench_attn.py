"""Micro-benchmark of the libmdv2 attention on the STDiT3-XL shapes (CUDA events)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from magicdrive_v2_b200 import ops

def bench(n_seq, Lq, Lk, H=16, D=72, kind="self", iters=5):
    bf = torch.bfloat16
    C = H * D
    qkv = torch.randn(n_seq * Lq, 3 * C, device="cuda", dtype=bf)
    q = qkv[:, :C]
    kv_map = None
    n_sets = 1
    if kind == "cross":
        kvb = torch.randn(n_seq * Lk, 2 * C, device="cuda", dtype=bf)
        k, v = kvb[:, :C], kvb[:, C:]
    else:
        k, v = qkv[:, C:2 * C], qkv[:, 2 * C:]
        if kind == "xview":
            kv_map = torch.stack([(torch.arange(n_seq) + 1) % n_seq, (torch.arange(n_seq) + n_seq - 1) % n_seq], 1).int().contiguous().cuda()
            n_sets = 2
    out = torch.empty(n_seq * Lq, C, device="cuda", dtype=bf)
    for _ in range(2):
        ops.attention(q, k, v, out, n_seq, Lq, Lk, H, D, D ** -0.5, kv_map=kv_map)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.attention(q, k, v, out, n_seq, Lq, Lk, H, D, D ** -0.5, kv_map=kv_map)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    fl = 4.0 * n_seq * Lq * Lk * H * D * n_sets
    print(f"{kind:6s} n_seq={n_seq:3d} Lq={Lq:5d} Lk={Lk:5d} D={D}: {ms:8.3f} ms  {fl / ms / 1e9:7.1f} TF/s (algorithmic, D un-padded)", flush=True)

if __name__ == "__main__":
    bench(60, 5600, 5600, iters=int(os.environ.get("ITERS", "5")))
    if "--quick" not in sys.argv:
        bench(60, 5600, 5600, kind="xview")
        bench(60, 5600, 62, kind="cross")
        bench(60, 400, 400)
        bench(60, 1600, 1600)
