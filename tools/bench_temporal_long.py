"""Micro-benchmark of the long-clip temporal attention (T = 17: 65 pixel frames, 424x800 -> S = 1600) at the
configs[3] size; the ncu capture of the same command is profiles/ncu_temporal_long_r01.txt."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from magicdrive_v2_b200 import ops  # noqa: E402
from tools.bench_elem import timeit  # noqa: E402

if __name__ == "__main__":
    bf = torch.bfloat16
    T = int(sys.argv[1]) if len(sys.argv) > 1 else 17
    b, NC, S, C, H, D = 2, 6, 1600, 1152, 16, 72
    N = b * NC * T * S
    qkv = torch.randn(N, 3 * C, device="cuda", dtype=bf)
    y = torch.empty(N, C, device="cuda", dtype=bf)
    cos = torch.randn(T, D, device="cuda")
    sin = torch.randn(T, D, device="cuda")
    ms = timeit(lambda: ops.temporal_attention(qkv, cos, sin, y, b * NC, T, S, H, D, D ** -0.5))
    print(f"temporal_attention T={T} {ms:7.3f} ms  {4 * N * C * 2 / ms / 1e6:7.1f} GB/s", flush=True)
