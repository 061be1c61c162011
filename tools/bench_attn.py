"""Micro-benchmark of the libmdv2 attention on the STDiT3-XL shapes (CUDA events), next to the same-box library
yard-sticks SURVEY.md §2.3 names: flash-attn (the call the reference makes, blocks.py:481-487) and cuDNN SDPA
(F.scaled_dot_product_attention).  The libraries are measured here only — the product never imports them.

  python tools/bench_attn.py             # ours on every shape + accuracy
  python tools/bench_attn.py --yardstick # ours / FA2 / cuDNN SDPA / (torch default SDPA) per shape, bf16, D = 72
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from magicdrive_v2_b200 import ops

BF = torch.bfloat16


def _time(fn, iters, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def _inputs(n_seq, Lq, Lk, H, D, kind):
    C = H * D
    qkv = torch.randn(n_seq * Lq, 3 * C, device="cuda", dtype=BF)
    q = qkv[:, :C]
    kv_map, n_sets = None, 1
    if kind == "cross":
        kvb = torch.randn(n_seq * Lk, 2 * C, device="cuda", dtype=BF)
        k, v = kvb[:, :C], kvb[:, C:]
    else:
        k, v = qkv[:, C:2 * C], qkv[:, 2 * C:]
        if kind == "xview":
            kv_map = torch.stack([(torch.arange(n_seq) + 1) % n_seq, (torch.arange(n_seq) + n_seq - 1) % n_seq],
                                 1).int().contiguous().cuda()
            n_sets = 2
    return q, k, v, kv_map, n_sets


def bench(n_seq, Lq, Lk, H=16, D=72, kind="self", iters=5):
    q, k, v, kv_map, n_sets = _inputs(n_seq, Lq, Lk, H, D, kind)
    out = torch.empty(n_seq * Lq, H * D, device="cuda", dtype=BF)
    ms = _time(lambda: ops.attention(q, k, v, out, n_seq, Lq, Lk, H, D, D ** -0.5, kv_map=kv_map), iters)
    fl = 4.0 * n_seq * Lq * Lk * H * D * n_sets
    print(f"{kind:6s} n_seq={n_seq:3d} Lq={Lq:5d} Lk={Lk:5d} D={D}: {ms:8.3f} ms  {fl / ms / 1e9:7.1f} TF/s "
          f"(algorithmic, D un-padded)", flush=True)
    return ms


def accuracy(n_seq=2, L=1408, H=16, D=72, Lk=None):
    """rel. L2 error against an fp32 softmax(QK^T)V of the same bf16 inputs."""
    C = H * D
    Lk = Lk or L
    g = torch.Generator(device="cuda").manual_seed(1)
    qb = torch.randn(n_seq * L, C, device="cuda", dtype=BF, generator=g)
    kvb = torch.randn(n_seq * Lk, 2 * C, device="cuda", dtype=BF, generator=g)
    out = torch.empty(n_seq * L, C, device="cuda", dtype=BF)
    ops.attention(qb, kvb[:, :C], kvb[:, C:], out, n_seq, L, Lk, H, D, D ** -0.5)
    q = qb.float().view(n_seq, L, H, D).transpose(1, 2)
    k, v = (kvb[:, i * C:(i + 1) * C].float().view(n_seq, Lk, H, D).transpose(1, 2) for i in range(2))
    ref = torch.softmax(q @ k.transpose(-1, -2) * D ** -0.5, dim=-1) @ v
    ref = ref.transpose(1, 2).reshape(n_seq * L, C)
    err = ((out.float() - ref).norm() / ref.norm()).item()
    print(f"accuracy vs fp32 reference (n_seq={n_seq}, Lq={L}, Lk={Lk}): rel L2 = {err:.3e}", flush=True)
    return err


def yardstick(n_seq, Lq, Lk, H=16, D=72, iters=5):
    """ours / flash-attn 2 / cuDNN SDPA / torch's default SDPA on one shape; the libraries get their own preferred
    layouts ([n, L, H, D] contiguous for flash-attn, [n, H, L, D] for SDPA), so no copy is inside their timed region."""
    import torch.nn.functional as F
    kind = "cross" if Lq != Lk else "self"
    q, k, v, _, _ = _inputs(n_seq, Lq, Lk, H, D, kind)
    out = torch.empty(n_seq * Lq, H * D, device="cuda", dtype=BF)
    fl = 4.0 * n_seq * Lq * Lk * H * D
    row = {}
    row["mdv2"] = _time(lambda: ops.attention(q, k, v, out, n_seq, Lq, Lk, H, D, D ** -0.5), iters)
    q4 = q.reshape(n_seq, Lq, H, D).contiguous()
    k4 = k.reshape(n_seq, Lk, H, D).contiguous()
    v4 = v.reshape(n_seq, Lk, H, D).contiguous()
    try:
        from flash_attn import flash_attn_func
        row["flash_attn2"] = _time(lambda: flash_attn_func(q4, k4, v4, softmax_scale=D ** -0.5), iters)
        ref = flash_attn_func(q4, k4, v4, softmax_scale=D ** -0.5).reshape(n_seq * Lq, H * D)
        row["_maxdiff_vs_fa2"] = (ref.float() - out.float()).abs().max().item()
    except Exception as e:  # noqa: BLE001
        row["flash_attn2"] = f"unavailable: {type(e).__name__}: {str(e)[:80]}"
    qh, kh, vh = (t.transpose(1, 2).contiguous() for t in (q4, k4, v4))
    try:
        from torch.nn.attention import SDPBackend, sdpa_kernel
        with sdpa_kernel(SDPBackend.CUDNN_ATTENTION):
            row["cudnn_sdpa"] = _time(lambda: F.scaled_dot_product_attention(qh, kh, vh, scale=D ** -0.5), iters)
    except Exception as e:  # noqa: BLE001
        row["cudnn_sdpa"] = f"unavailable: {type(e).__name__}: {str(e)[:80]}"
    try:
        row["torch_sdpa_default"] = _time(lambda: F.scaled_dot_product_attention(qh, kh, vh, scale=D ** -0.5), iters)
    except Exception as e:  # noqa: BLE001
        row["torch_sdpa_default"] = f"unavailable: {type(e).__name__}: {str(e)[:80]}"
    cells = []
    for name in ("mdv2", "flash_attn2", "cudnn_sdpa", "torch_sdpa_default"):
        r = row[name]
        cells.append(f"{name} {r:8.3f} ms {fl / r / 1e9:7.1f} TF/s" if isinstance(r, float) else f"{name} {r}")
    extra = f"  max|ours-fa2| {row['_maxdiff_vs_fa2']:.3e}" if "_maxdiff_vs_fa2" in row else ""
    print(f"n_seq={n_seq:3d} Lq={Lq:5d} Lk={Lk:5d} H={H} D={D} bf16: " + " | ".join(cells) + extra, flush=True)


if __name__ == "__main__":
    if "--yardstick" in sys.argv:
        print(torch.cuda.get_device_name(0), "torch", torch.__version__, flush=True)
        for (n, lq, lk) in ((60, 5600, 5600), (204, 1600, 1600), (60, 1400, 1400), (60, 400, 400), (60, 5600, 62)):
            yardstick(n, lq, lk)
        sys.exit(0)
    accuracy()
    accuracy(L=1600)
    accuracy(L=333, Lk=62)
    bench(60, 5600, 5600, iters=int(os.environ.get("ITERS", "5")))
    if "--quick" not in sys.argv:
        bench(60, 5600, 5600, kind="xview")
        bench(60, 5600, 62, kind="cross")
        bench(60, 400, 400)
        bench(204, 1600, 1600)
        bench(60, 1400, 1400)
