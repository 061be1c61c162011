"""ctypes binding of ``libmdv2.so`` — the only execution path of the STDiT3 hot loop.

Every function takes torch CUDA tensors, passes raw ``data_ptr()``s plus sizes through the C ABI
declared in ``include/mdv2.h`` and launches on torch's current stream.  There is deliberately no
CPU or eager fallback: importing this module without the built library, or calling an op with a
non-CUDA tensor, raises.
"""
from __future__ import annotations

import ctypes
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MDV2_LIB_FILE") or os.path.join(_HERE, "libmdv2.so")   # (MDV2_LIB_FILE: A/B builds in tools/)

BACKEND_AUTO, BACKEND_SIMT, BACKEND_TCGEN05 = 0, 1, 2
_BACKENDS = {None: 0, "auto": 0, "simt": 1, "tcgen05": 2}

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -m magicdrive_v2_b200.build` "
        "(or __graft_entry__.build()); the STDiT3 drop-in has no fallback path.")

_lib = ctypes.CDLL(LIB_PATH)
_vp, _i64, _i32, _f32 = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_float

_lib.mdv2_version.restype = _i32
_lib.mdv2_last_error.restype = ctypes.c_char_p
_lib.mdv2_last_kernel.restype = ctypes.c_char_p
_lib.mdv2_ln_modulate.argtypes = [_vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _i32, _i32, _f32, _i32, _vp]
_lib.mdv2_adaln_tables.argtypes = [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _vp]
_lib.mdv2_gemm.argtypes = [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _i64, _i32, _i32, _i32, _f32, _vp, _i32, _i32,
                           _vp, _i64, _i64, _i64, _i32, _i32, _i32, _vp, _vp, _i32, _i32, _i32, _f32, _i32, _i32, _vp]
_lib.mdv2_attention.argtypes = [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _f32,
                                _vp, _i32, _i32, _i32, _vp]
_lib.mdv2_temporal_attention.argtypes = [_vp, _i64, _vp, _vp, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _f32,
                                         _i32, _vp]
_lib.mdv2_patchify.argtypes = [_vp, _vp] + [_i32] * 11 + [_vp]
_lib.mdv2_unpatchify.argtypes = [_vp, _vp] + [_i32] * 10 + [_vp]
_lib.mdv2_swap01.argtypes = [_vp, _vp, _i64, _i64, _i64, _vp]
_lib.mdv2_scatter_runs.argtypes = [_vp, ctypes.POINTER(_vp), _i32, _i32, ctypes.POINTER(_i64), ctypes.POINTER(_i64),
                                   ctypes.POINTER(_i64), _i64, _i64, _vp]
_lib.mdv2_cfg_euler_step.argtypes = [_vp, _vp, _i32, _vp, _f32, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _i32, _i64, _vp]
_lib.mdv2_gemm_qkv_scatter.argtypes = [_vp, _i64, _vp, _i64, _vp, _i64, _i32, _i32, _vp, _vp, _i32, _f32,
                                       ctypes.POINTER(_vp), _i32, _i32, _i32, _vp]
_lib.mdv2_attention_scatter.argtypes = [_vp, _i64, _vp, _i64, _vp, _i64, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _f32,
                                        _vp, _i32, ctypes.POINTER(_vp), _i32, _i32, _i64, _i32, _vp]
_lib.mdv2_attention_scatter_static.argtypes = [_vp, _i64, _vp, _i64, _vp, _i64, _i32, _i32, _i32, _i32, _i32, _f32, _vp,
                                               _i32, ctypes.POINTER(_vp), _i32, _i32, _i64, _i32, _vp]
_lib.mdv2_static_scratch_bytes.restype = _i64
for _f in ("mdv2_adaln_tables", "mdv2_attention_scatter_static", "mdv2_attention_scatter", "mdv2_gemm_qkv_scatter", "mdv2_cfg_euler_step", "mdv2_ln_modulate", "mdv2_gemm", "mdv2_attention", "mdv2_temporal_attention", "mdv2_patchify",
           "mdv2_unpatchify", "mdv2_swap01", "mdv2_scatter_runs"):
    getattr(_lib, _f).restype = _i32

EXPORTS = ("mdv2_version", "mdv2_last_error", "mdv2_last_kernel", "mdv2_ln_modulate", "mdv2_gemm", "mdv2_attention",
           "mdv2_temporal_attention", "mdv2_patchify", "mdv2_unpatchify", "mdv2_swap01", "mdv2_scatter_runs", "mdv2_cfg_euler_step", "mdv2_gemm_qkv_scatter",
           "mdv2_attention_scatter", "mdv2_attention_scatter_static", "mdv2_static_scratch_bytes", "mdv2_adaln_tables")

# kernel-launch counter (bench.py reports it as gpu_launches); +1 per C-ABI call that launches
launches = 0
default_backend = _BACKENDS[os.environ.get("MDV2_BACKEND", "auto")]


def version() -> int:
    return _lib.mdv2_version()


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return 0
    if t.dtype == torch.bfloat16:
        return 1
    raise TypeError(f"libmdv2 supports fp32 and bf16 tensors, got {t.dtype}")


def _ptr(t, dtype=None):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("libmdv2 ops need CUDA tensors (there is no CPU fallback)")
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"expected {dtype}, got {t.dtype}")
    return t.data_ptr()


def _rows2d(t: torch.Tensor):
    """(ptr, leading dimension) of a 2-D view whose last dim is contiguous."""
    assert t.dim() == 2 and (t.stride(1) == 1 or t.shape[1] == 1), "need a row-major 2-D view"
    return _ptr(t), t.stride(0)


def _stream():
    return torch.cuda.current_stream().cuda_stream


# per-call device timing (bench.py's kernel breakdown): a list of
# (kernel family, algorithmic flops, algorithmic bytes, start event, end event) while enabled
_prof = None


def profile_start():
    global _prof
    _prof = []


def profile_stop():
    """-> {kernel family: dict(calls, ms, flops, bytes)}; synchronises."""
    global _prof
    rec, _prof = _prof, None
    torch.cuda.synchronize()
    out = {}
    for kind, fl, by, e0, e1 in rec or []:
        d = out.setdefault(kind, dict(calls=0, ms=0.0, flops=0.0, bytes=0.0))
        d["calls"] += 1
        d["ms"] += e0.elapsed_time(e1)
        d["flops"] += fl
        d["bytes"] += by
    return out


def _t0():
    if _prof is None:
        return None
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


def _t1(e0, tag, flops, nbytes):
    if e0 is None:
        return
    e1 = torch.cuda.Event(enable_timing=True)
    e1.record()
    kind = _lib.mdv2_last_kernel().decode()
    _prof.append((kind + tag, float(flops), float(nbytes), e0, e1))


def _check(rc: int, name: str):
    global launches
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {_lib.mdv2_last_error().decode()}")
    launches += 1


def adaln_tables(tables, tsel, out=None):
    """fp32 ``[n, BT, K, C]`` = tables ``[n, K, C]`` + tsel ``[BT, K, C or 1]``, the sum rounded to the model dtype."""
    n, K, C = tables.shape
    BT, Ct = tsel.shape[0], tsel.shape[-1]
    assert tables.is_contiguous() and tsel.is_contiguous() and tsel.shape[1] == K and tsel.dtype == tables.dtype
    if out is None:
        out = torch.empty(n, BT, K, C, dtype=torch.float32, device=tables.device)
    e0 = _t0()
    _check(_lib.mdv2_adaln_tables(_ptr(tables), _ptr(tsel), _ptr(out, torch.float32), n, BT, K, C, Ct, _dt(tables),
                                  _stream()), "mdv2_adaln_tables")
    _t1(e0, "", 0.0, out.numel() * 4 + tables.numel() * tables.element_size())
    return out


def ln_modulate(x, out, mod, k_shift, k_scale, geo, eps=1e-6):
    T, S, NC = geo
    N, C = x.shape
    assert x.is_contiguous() and out.is_contiguous() and mod.is_contiguous() and mod.dtype == torch.float32
    assert mod.shape[-1] == C and out.shape == x.shape and out.dtype == x.dtype
    e0 = _t0()
    _check(_lib.mdv2_ln_modulate(_ptr(x), _ptr(out), _ptr(mod), N, C, mod.shape[-2], k_shift, k_scale, T, S, NC,
                                 eps, _dt(x), _stream()), "mdv2_ln_modulate")
    _t1(e0, "", 0, 2 * N * C * x.element_size())
    return out


def gemm(a, w, bias, out, geo=None, act=None, bias_scale=1.0, gate=None, res=None, res_map=None, rms=None,
         backend=None):
    M, K = a.shape
    N = w.shape[0]
    assert w.shape[1] == K and out.shape == (M, N) and a.dtype == w.dtype == out.dtype
    pa, lda = _rows2d(a)
    pw, ldw = _rows2d(w)
    po, ldo = _rows2d(out)
    T, S, NC = geo if geo is not None else (1, 1, 1)
    gm, modK, kg = None, 0, 0
    if gate is not None:
        mod, kg = gate
        assert mod.dtype == torch.float32 and mod.is_contiguous() and mod.shape[-1] == N and geo is not None
        gm, modK = _ptr(mod), mod.shape[-2]
    pr, ldr, rmod, rdiv = None, 0, 0, 0
    if res is not None:
        assert res.dtype == a.dtype and res.shape[1] == N
        pr, ldr = _rows2d(res)
        if res_map is not None:
            rmod, rdiv = res_map
        else:
            assert res.shape[0] == M
    qw = kw = None
    hd = nq = nk = 0
    if rms is not None and rms[0] is not None:
        qw_t, kw_t, hd, nq, nk = rms
        qw, kw = _ptr(qw_t, a.dtype), _ptr(kw_t, a.dtype)
    be = default_backend if backend is None else _BACKENDS[backend]
    e0 = _t0()
    _check(_lib.mdv2_gemm(pa, lda, pw, ldw, _ptr(bias, a.dtype) if bias is not None else None, po, ldo, M, N, K,
                          1 if act == "gelu" else 0, float(bias_scale), gm, modK, kg, pr, ldr, rmod, rdiv, T, S, NC,
                          qw, kw, hd, nq, nk, 1e-6, _dt(a), be, _stream()), "mdv2_gemm")
    _t1(e0, "", 2.0 * M * N * K, (M * K + N * K + M * N * (2 if res is not None else 1)) * a.element_size())
    return out


def attention(q, k, v, out, n_seq, Lq, Lk, H, D, scale, kv_map=None, backend=None):
    pq, ldq = _rows2d(q)
    pk, ldk = _rows2d(k)
    pv, ldv = _rows2d(v)
    po, ldo = _rows2d(out)
    assert q.shape[0] == n_seq * Lq and q.shape[1] == H * D and out.shape == q.shape
    assert k.shape[1] == H * D and v.shape == k.shape and k.shape[0] % Lk == 0
    n_sets, pm = 1, None
    if kv_map is not None:
        assert kv_map.dtype == torch.int32 and kv_map.is_contiguous() and kv_map.shape[0] == n_seq
        n_sets, pm = kv_map.shape[1], _ptr(kv_map)
    be = default_backend if backend is None else _BACKENDS[backend]
    e0 = _t0()
    _check(_lib.mdv2_attention(pq, ldq, pk, ldk, pv, ldv, po, ldo, n_seq, Lq, Lk, H, D, float(scale), pm, n_sets,
                               _dt(q), be, _stream()), "mdv2_attention")
    # the spatial self-attention and the cross-view attention are ONE kernel (K/V sets swept back to back): one row
    tag = "[cross]" if Lk <= 512 and Lq != Lk else "[spatial self + cross-view]"
    _t1(e0, tag, 4.0 * n_seq * Lq * Lk * H * D * n_sets,
        (2 * q.shape[0] + 2 * n_sets * n_seq * Lk) * H * D * q.element_size())
    return out


def temporal_attention(qkv, cos, sin, out, B, T, S, H, D, scale):
    pq, ld = _rows2d(qkv)
    po, ldo = _rows2d(out)
    assert qkv.shape == (B * T * S, 3 * H * D) and out.shape == (B * T * S, H * D)
    assert cos.dtype == torch.float32 and cos.is_contiguous() and cos.shape == (T, D) and sin.shape == (T, D)
    e0 = _t0()
    _check(_lib.mdv2_temporal_attention(pq, ld, _ptr(cos), _ptr(sin), po, ldo, B, T, S, H, D, float(scale),
                                        _dt(qkv), _stream()), "mdv2_temporal_attention")
    _t1(e0, "", 4.0 * B * S * T * T * H * D, 4 * B * T * S * H * D * qkv.element_size())
    return out


def patchify(x5d, out, NCi, Hp, Wp, s0, Sl):
    Bx, CN, T, Hx, Wx = x5d.shape
    Cin = CN // NCi
    assert x5d.is_contiguous() and out.is_contiguous() and out.dtype == x5d.dtype
    assert out.shape == (Bx * NCi * T * Sl, Cin * 4)
    _check(_lib.mdv2_patchify(_ptr(x5d), _ptr(out), Bx, Cin, NCi, T, Hx, Wx, Hp, Wp, s0, Sl, _dt(x5d), _stream()),
           "mdv2_patchify")
    return out


def unpatchify(yo, out, NC, T, Hp, Wp):
    b, CN, Tx, Hx, Wx = out.shape
    Co = CN // NC
    assert yo.is_contiguous() and out.is_contiguous() and out.dtype == torch.float32
    assert yo.shape == (b * NC * T * Hp * Wp, 4 * Co)
    _check(_lib.mdv2_unpatchify(_ptr(yo), _ptr(out), b, NC, Co, T, Hp, Wp, Tx, Hx, Wx, _dt(yo), _stream()),
           "mdv2_unpatchify")
    return out


def swap01(src, dst, A, B, inner):
    """dst[b][a][:] = src[a][b][:]; ``inner`` elements per contiguous run (both tensors contiguous, same dtype)."""
    assert src.is_contiguous() and dst.is_contiguous() and src.dtype == dst.dtype
    assert src.numel() == dst.numel() == A * B * inner
    e0 = _t0()
    _check(_lib.mdv2_swap01(_ptr(src), _ptr(dst), A, B, inner * src.element_size(), _stream()), "mdv2_swap01")
    _t1(e0, "", 0, 2 * src.numel() * src.element_size())
    return dst


def scatter_runs(src, dst_ptrs, peer_dim, sizes, src_strides, dst_strides, run, dst_base=0):
    """Strided scatter of contiguous runs of ``run`` elements of ``src`` into the buffers ``dst_ptrs`` (raw device
    addresses, e.g. the NVLink-mapped peers of a symmetric allocation); index ``peer_dim`` selects the destination.
    Strides, ``run`` and ``dst_base`` are in elements of ``src``'s dtype."""
    es = src.element_size()
    n = len(dst_ptrs)
    arr = (_vp * n)(*[_vp(int(q)) for q in dst_ptrs])
    I4 = _i64 * 4
    e0 = _t0()
    _check(_lib.mdv2_scatter_runs(_ptr(src), arr, n, peer_dim, I4(*sizes), I4(*[x * es for x in src_strides]),
                                  I4(*[x * es for x in dst_strides]), run * es, dst_base * es, _stream()),
           "mdv2_scatter_runs")
    nb = run * es
    for x in sizes:
        nb *= x
    _t1(e0, "", 0, 2 * nb)


def cfg_euler_step(pred, z, dt, guidance, x0=None, keep=None, z_in=None):
    """Fused CFG combine + Euler update (+ frame-mask restore, + the next forward's doubled, cast input).

    pred fp32 [2n, C, T, H, W] (cond half first), z [n, C, T, H, W] fp32 / bf16, dt fp32 [n] on the device,
    keep bool/uint8 [n, T] with x0 (z's dtype) or both None, z_in optional [2n, C, T, H, W] fp32 / bf16.
    Returns the fp32 z' (written in place when z is fp32)."""
    n, C, T = z.shape[0], z.shape[1], z.shape[2]
    hw = z.shape[3] * z.shape[4]
    assert pred.dtype == torch.float32 and pred.is_contiguous() and pred.shape[0] == 2 * n
    assert tuple(pred.shape[1:]) == tuple(z.shape[1:]) and z.is_contiguous()
    assert dt.dtype == torch.float32 and dt.numel() == n and dt.is_contiguous()
    z_out = z if z.dtype == torch.float32 else torch.empty_like(z, dtype=torch.float32)
    if keep is not None:
        assert x0 is not None and x0.dtype == z.dtype and x0.shape == z.shape and x0.is_contiguous()
        keep = keep.to(torch.uint8).contiguous()
        assert tuple(keep.shape) == (n, T)
    if z_in is not None:
        assert z_in.is_contiguous() and z_in.shape[0] == 2 * n and tuple(z_in.shape[1:]) == tuple(z.shape[1:])
    e0 = _t0()
    _check(_lib.mdv2_cfg_euler_step(_ptr(pred), _ptr(z), _dt(z), _ptr(dt), float(guidance),
                                    _ptr(x0) if x0 is not None else None, _ptr(keep) if keep is not None else None,
                                    _ptr(z_out), _ptr(z_in) if z_in is not None else None,
                                    _dt(z_in) if z_in is not None else 0, n, C, T, hw, _stream()),
           "mdv2_cfg_euler_step")
    _t1(e0, "", 0, pred.numel() * 4 + z.numel() * (z.element_size() + 4)
        + (z_in.numel() * z_in.element_size() if z_in is not None else 0))
    return z_out


def gemm_qkv_scatter_supported(a, w, head_dim, p) -> bool:
    """The fused projection + scatter exists for the bf16 tcgen05 kernel with whole heads per rank."""
    N = w.shape[0]
    return (a.dtype == torch.bfloat16 and default_backend != BACKEND_SIMT and head_dim in (64, 72) and N % 3 == 0
            and (N // 3) % p == 0 and ((N // 3) // p) % head_dim == 0 and a.shape[1] % 64 == 0 and p <= 8)


def gemm_qkv_scatter(a, w, bias, rms, dst_ptrs, rows_per_seq, rank):
    """qkv = RMSNorm_qk(a @ w.T + bias), head group g of q | k | v stored into ``dst_ptrs[g]`` (raw device addresses
    of the peers' ``[seq, p·rows_per_seq, 3, W/p]`` buffers) at sequence rows ``rank·rows_per_seq + r``."""
    wq, wk, D, n_q, n_k = rms
    M, K = a.shape
    N = w.shape[0]
    assert n_q == n_k == N // 3 and w.shape[1] == K and M % rows_per_seq == 0
    pa, lda = _rows2d(a)
    pw, ldw = _rows2d(w)
    n = len(dst_ptrs)
    arr = (_vp * n)(*[_vp(int(q)) for q in dst_ptrs])
    e0 = _t0()
    _check(_lib.mdv2_gemm_qkv_scatter(pa, lda, pw, ldw, _ptr(bias, a.dtype) if bias is not None else None, M, N, K,
                                      _ptr(wq, a.dtype), _ptr(wk, a.dtype), D, 1e-6, arr, n, rows_per_seq, rank,
                                      _stream()), "mdv2_gemm_qkv_scatter")
    _t1(e0, "", 2.0 * M * N * K, (M * K + N * K + M * N) * a.element_size())


def attention_scatter_supported(q, D, p) -> bool:
    return q.dtype == torch.bfloat16 and default_backend != BACKEND_SIMT and D in (64, 72) and p <= 8


def attention_scatter(q, k, v, partial, n_seq, Lq, Lk, H, D, scale, dst_ptrs, rows_per_rank, dst_ld, dst_col,
                      kv_map=None):
    """``attention`` whose final output rows go straight into ``dst_ptrs[row // rows_per_rank]`` (the peers' ``[N_l, C]``
    buffers, this rank's head group at column ``dst_col``); ``partial`` holds the sum of earlier K/V sets."""
    pq, ldq = _rows2d(q)
    pk, ldk = _rows2d(k)
    pv, ldv = _rows2d(v)
    n_sets = 1 if kv_map is None else kv_map.shape[1]
    pp, ldp = _rows2d(partial)
    pm = _ptr(kv_map, torch.int32) if kv_map is not None else None
    n = len(dst_ptrs)
    arr = (_vp * n)(*[_vp(int(a)) for a in dst_ptrs])
    e0 = _t0()
    _check(_lib.mdv2_attention_scatter(pq, ldq, pk, ldk, pv, ldv, pp, ldp, n_seq, Lq, Lk, H, D, float(scale), pm,
                                       n_sets, arr, n, rows_per_rank, dst_ld, dst_col, _stream()),
           "mdv2_attention_scatter")
    _t1(e0, "[spatial self + cross-view]", 4.0 * n_seq * Lq * Lk * H * D * n_sets,
        (2 * q.shape[0] + 2 * n_sets * n_seq * Lk) * H * D * q.element_size())


def static_scratch_bytes() -> int:
    return int(_lib.mdv2_static_scratch_bytes())


def attention_scatter_static(q, k, v, n_seq, Lq, Lk, H, D, scale, dst_ptrs, rows_per_rank, dst_ld, dst_col, kv_map=None):
    """``attention_scatter`` without a caller-owned ``partial`` buffer: the sum of earlier K/V sets lives in the
    library's static device array (``static_scratch_bytes()``); allocates nothing.  Calls that overlap on the GPU
    race on that array."""
    pq, ldq = _rows2d(q)
    pk, ldk = _rows2d(k)
    pv, ldv = _rows2d(v)
    n_sets = 1 if kv_map is None else kv_map.shape[1]
    pm = _ptr(kv_map, torch.int32) if kv_map is not None else None
    n = len(dst_ptrs)
    arr = (_vp * n)(*[_vp(int(a)) for a in dst_ptrs])
    _check(_lib.mdv2_attention_scatter_static(pq, ldq, pk, ldk, pv, ldv, n_seq, Lq, Lk, H, D, float(scale), pm, n_sets,
                                              arr, n, rows_per_rank, dst_ld, dst_col, _stream()),
           "mdv2_attention_scatter_static")
