"""Micro-benchmark of the HBM-bound libmdv2 kernels at the cfg3 size (336 000 tokens)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from magicdrive_v2_b200 import ops

def timeit(fn, iters=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

if __name__ == "__main__":
    bf = torch.bfloat16
    b, NC, T, S, C, H, D = 2, 6, 5, 5600, 1152, 16, 72
    N = b * NC * T * S
    x = torch.randn(N, C, device="cuda", dtype=bf)
    y = torch.empty_like(x)
    mod = torch.randn(b * T, 6, C, device="cuda", dtype=torch.float32) * 0.1
    ms = timeit(lambda: ops.ln_modulate(x, y, mod, 0, 1, (T, S, NC)))
    print(f"ln_modulate        {ms:7.3f} ms  {2 * N * C * 2 / ms / 1e6:7.1f} GB/s", flush=True)
    qkv = torch.randn(N, 3 * C, device="cuda", dtype=bf)
    cos = torch.randn(T, D, device="cuda"); sin = torch.randn(T, D, device="cuda")
    ms = timeit(lambda: ops.temporal_attention(qkv, cos, sin, y, b * NC, T, S, H, D, D ** -0.5))
    print(f"temporal_attention {ms:7.3f} ms  {4 * N * C * 2 / ms / 1e6:7.1f} GB/s", flush=True)
