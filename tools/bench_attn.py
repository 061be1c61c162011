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

def accuracy(n_seq=2, L=1408, H=16, D=72):
    """rel. L2 error against an fp32 softmax(QK^T)V of the same bf16 inputs."""
    bf = torch.bfloat16
    C = H * D
    g = torch.Generator(device="cuda").manual_seed(1)
    qkv = torch.randn(n_seq * L, 3 * C, device="cuda", dtype=bf, generator=g)
    out = torch.empty(n_seq * L, C, device="cuda", dtype=bf)
    ops.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], out, n_seq, L, L, H, D, D ** -0.5)
    q, k, v = (qkv[:, i * C:(i + 1) * C].float().view(n_seq, L, H, D).transpose(1, 2) for i in range(3))
    ref = torch.softmax(q @ k.transpose(-1, -2) * D ** -0.5, dim=-1) @ v
    ref = ref.transpose(1, 2).reshape(n_seq * L, C)
    err = ((out.float() - ref).norm() / ref.norm()).item()
    print(f"accuracy vs fp32 reference (n_seq={n_seq}, L={L}): rel L2 = {err:.3e}", flush=True)


if __name__ == "__main__":
    accuracy()
    bench(60, 5600, 5600, iters=int(os.environ.get("ITERS", "5")))
    if "--quick" not in sys.argv:
        bench(60, 5600, 5600, kind="xview")
        bench(60, 5600, 62, kind="cross")
        bench(60, 400, 400)
        bench(60, 1600, 1600)
