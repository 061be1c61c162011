"""fc2-shaped GEMM (M=336000, N=1152, K=4608, bias + residual) for ncu: DRAM traffic vs the number of CTA pairs
(MDV2_GEMM_PAIRS).  With 74 pairs and 6 column tiles per row block the groups of pairs that share an A row block
re-shuffle every round; 72 pairs keep them fixed."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from magicdrive_v2_b200 import ops
M, N, K = 336000, 1152, 4608
bf = torch.bfloat16
a = torch.randn(M, K, device="cuda", dtype=bf)
w = torch.randn(N, K, device="cuda", dtype=bf) * K ** -0.5
b = torch.randn(N, device="cuda", dtype=bf)
x = torch.randn(M, N, device="cuda", dtype=bf)
for _ in range(3):
    ops.gemm(a, w, b, x, res=x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    ops.gemm(a, w, b, x, res=x)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 10
print(f"pairs={os.environ.get('MDV2_GEMM_PAIRS', 'all')}: {ms:.3f} ms  {2.0 * M * N * K / ms / 1e9:.1f} TF/s")
