"""Micro-benchmark of the libmdv2 GEMM on the STDiT3-XL shapes (CUDA events, L2-cold rotation)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from magicdrive_v2_b200 import ops

def bench(M, N, K, iters=20, **kw):
    bf = torch.bfloat16
    nbuf = max(2, int(300e6 // (M * K * 2 + M * N * 2)) + 1)
    a = [torch.randn(M, K, device="cuda", dtype=bf) for _ in range(nbuf)]
    o = [torch.empty(M, N, device="cuda", dtype=bf) for _ in range(nbuf)]
    w = torch.randn(N, K, device="cuda", dtype=bf) * K ** -0.5
    b = torch.randn(N, device="cuda", dtype=bf)
    for i in range(3):
        ops.gemm(a[i % nbuf], w, b, o[i % nbuf], **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        ops.gemm(a[i % nbuf], w, b, o[i % nbuf], **kw)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    tf = 2.0 * M * N * K / ms / 1e9
    # cuBLAS yard-stick
    for i in range(3): torch.matmul(a[i % nbuf], w.t())
    e0.record()
    for i in range(iters): torch.matmul(a[i % nbuf], w.t())
    e1.record(); torch.cuda.synchronize()
    ms2 = e0.elapsed_time(e1) / iters
    print(f"M={M:7d} N={N:5d} K={K:5d}  mdv2 {ms:8.3f} ms {tf:7.1f} TF   cublas {ms2:8.3f} ms {2.0*M*N*K/ms2/1e9:7.1f} TF", flush=True)

def sustained(M, N, K, secs=1.5, **kw):
    """Back-to-back launches for `secs` seconds (power-capped clocks), ours vs cuBLAS."""
    bf = torch.bfloat16
    a = torch.randn(M, K, device="cuda", dtype=bf)
    o = torch.empty(M, N, device="cuda", dtype=bf)
    w = torch.randn(N, K, device="cuda", dtype=bf) * K ** -0.5
    b = torch.randn(N, device="cuda", dtype=bf)
    res = {}
    for name, fn in (("mdv2", lambda: ops.gemm(a, w, b, o, **kw)), ("cublas", lambda: torch.matmul(a, w.t(), out=o))):
        fn(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        n = max(10, int(secs * 1e3 / max(e0.elapsed_time(e1), 1e-3)))
        e0.record()
        for _ in range(n):
            fn()
        e1.record(); torch.cuda.synchronize()
        res[name] = 2.0 * M * N * K * n / e0.elapsed_time(e1) / 1e9
    print(f"sustained M={M} N={N} K={K}: mdv2 {res['mdv2']:7.1f} TF  cublas {res['cublas']:7.1f} TF", flush=True)


def model_cases(M=84000, iters=20):
    """The five GEMM flavours of a block with their in-model epilogues (T=5, S=M/60... geometry of cfg3)."""
    bf = torch.bfloat16
    C = 1152
    S = M // 60
    geo = (5, S, 6)
    mod = torch.randn(10, 6, C, device="cuda", dtype=torch.float32) * 0.1
    qn = torch.ones(72, device="cuda", dtype=bf)
    def run(name, N, K, **kw):
        a = torch.randn(M, K, device="cuda", dtype=bf)
        w = torch.randn(N, K, device="cuda", dtype=bf) * K ** -0.5
        b = torch.randn(N, device="cuda", dtype=bf)
        o = torch.randn(M, N, device="cuda", dtype=bf)
        if kw.pop("inplace_res", False):
            kw["res"] = o
        fn = lambda: ops.gemm(a, w, b, o, backend="tcgen05", **kw)
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): fn()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / iters
        print(f"{name:22s} M={M} N={N:5d} K={K:5d}: {ms:7.3f} ms {2.0 * M * N * K / ms / 1e9:7.1f} TF", flush=True)
    run("proj  bias+gate+res", C, C, geo=geo, gate=(mod, 2), inplace_res=True)
    run("q_lin bias", C, C)
    run("fc2   bias+gate+res", C, 4 * C, geo=geo, gate=(mod, 5), inplace_res=True)
    run("fc1   bias+gelu", 4 * C, C, act="gelu")
    run("qkv   bias+rmsnorm", 3 * C, C, rms=(qn, qn, 72, C, C))


if __name__ == "__main__":
    if "--model" in sys.argv:
        model_cases()
        sys.exit(0)
    if "--profile" in sys.argv:
        torch.cuda.profiler.start()
        model_cases(iters=1)
        torch.cuda.profiler.stop()
        sys.exit(0)
    if "--sustained" in sys.argv:
        for (N, K) in ((1152, 1152), (3456, 1152), (4608, 1152), (1152, 4608)):
            sustained(336000, N, K, backend="tcgen05")
        sys.exit(0)
    for M in (24000, 336000 // 4):
        for (N, K) in ((1152, 1152), (3456, 1152), (4608, 1152), (1152, 4608), (2304, 1152)):
            bench(M, N, K, backend="tcgen05")
    bench(24000, 4608, 1152, backend="tcgen05", act="gelu")
