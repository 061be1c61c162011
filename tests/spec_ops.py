"""TEST INFRASTRUCTURE — plain-PyTorch statement of every entry point of ``magicdrive_v2_b200.ops``.

Each function has the signature and the in-place semantics of the CUDA op it specifies and computes
in fp32 (result rounded once to the output dtype).  Used (a) on the GPU as the per-kernel checker and
(b) on the CPU, injected into the drop-in, to test the host-side logic (index maps, schedules,
sequence-parallel plumbing) without a GPU.  The product never imports this file.
"""
import torch
import torch.nn.functional as F


def _mod_rows(n_rows, geo, device):
    T, S, NC = geo
    r = torch.arange(n_rows, device=device)
    g = r // S
    return (g // T // NC) * T + g % T


def adaln_tables(tables, tsel, out=None):
    r = (tables[:, None] + tsel[None]).float()               # the add rounds to the model dtype, like the reference's
    if out is None:
        return r.contiguous()
    out.copy_(r)
    return out


def ln_modulate(x, out, mod, k_shift, k_scale, geo, eps=1e-6):
    xf = x.float()
    xn = F.layer_norm(xf, (xf.shape[-1],), eps=eps)
    rows = _mod_rows(x.shape[0], geo, x.device)
    out.copy_((xn * (1 + mod[rows, k_scale]) + mod[rows, k_shift]).to(out.dtype))
    return out


def gemm(a, w, bias, out, geo=None, act=None, bias_scale=1.0, gate=None, res=None, res_map=None,
         rms=None, backend=None):
    acc = a.float() @ w.float().t()
    if bias is not None:
        acc = acc + bias_scale * bias.float()
    if rms is not None and rms[0] is not None:
        qw, kw, D, nq, nk = rms
        M = acc.shape[0]

        def norm(cols, wgt):
            v = cols.reshape(M, -1, D)
            v = v * torch.rsqrt(v.pow(2).mean(-1, keepdim=True) + 1e-6)
            return (v * wgt.float()).reshape(M, -1)
        acc = torch.cat([norm(acc[:, :nq], qw), norm(acc[:, nq:nq + nk], kw), acc[:, nq + nk:]], dim=1)
    if act == "gelu":
        acc = F.gelu(acc, approximate="tanh")
    if gate is not None:
        mod, kg = gate
        acc = acc * mod[_mod_rows(acc.shape[0], geo, a.device), kg]
    if res is not None:
        if res_map is not None:
            m, div = res_map
            r = torch.arange(acc.shape[0], device=a.device)
            rr = (r // (m * div)) * m + r % m if div > 0 else r % m
            acc = acc + res.float()[rr]
        else:
            acc = acc + res.float()
    out.copy_(acc.to(out.dtype))
    return out


def attention(q, k, v, out, n_seq, Lq, Lk, H, D, scale, kv_map=None, backend=None):
    qf = q.float().reshape(n_seq, Lq, H, D).permute(0, 2, 1, 3)
    kf = k.float().reshape(-1, Lk, H, D).permute(0, 2, 1, 3)
    vf = v.float().reshape(-1, Lk, H, D).permute(0, 2, 1, 3)
    if kv_map is None:
        maps = [torch.arange(n_seq, device=q.device)]
    else:
        maps = [kv_map[:, j].long() for j in range(kv_map.shape[1])]
    acc = 0
    for idx in maps:
        p = ((qf * scale) @ kf[idx].transpose(-1, -2)).softmax(dim=-1)
        acc = acc + p @ vf[idx]
    out.copy_(acc.permute(0, 2, 1, 3).reshape(n_seq * Lq, H * D).to(out.dtype))
    return out


def temporal_attention(qkv, cos, sin, out, B, T, S, H, D, scale):
    C = H * D
    x = qkv.float().reshape(B, T, S, 3, H, D).permute(3, 0, 2, 4, 1, 5)        # [3, B, S, H, T, D]
    q, k, v = x[0], x[1], x[2]

    def rope(t):
        t1, t2 = t[..., 0::2], t[..., 1::2]
        rot = torch.stack((-t2, t1), dim=-1).flatten(-2)
        return t * cos + rot * sin
    q, k = rope(q), rope(k)
    p = ((q * scale) @ k.transpose(-1, -2)).softmax(dim=-1)
    o = (p @ v).permute(0, 3, 1, 2, 4)                                           # [B, T, S, H, D]
    out.copy_(o.reshape(B * T * S, C).to(out.dtype))
    return out


def patchify(x5d, out, NCi, Hp, Wp, s0, Sl):
    Bx, CN, T, Hx, Wx = x5d.shape
    Cin = CN // NCi
    x = x5d.reshape(Bx, Cin, NCi, T, Hx, Wx).permute(0, 2, 1, 3, 4, 5).reshape(Bx * NCi, Cin, T, Hx, Wx)
    x = F.pad(x, (0, 2 * Wp - Wx, 0, 2 * Hp - Hx))
    x = x.reshape(Bx * NCi, Cin, T, Hp, 2, Wp, 2).permute(0, 2, 3, 5, 1, 4, 6)   # [β,T,Hp,Wp,Cin,2,2]
    x = x.reshape(Bx * NCi, T, Hp * Wp, Cin * 4)[:, :, s0:s0 + Sl]
    out.copy_(x.reshape(-1, Cin * 4))
    return out


def unpatchify(yo, out, NC, T, Hp, Wp):
    b, CN, Tx, Hx, Wx = out.shape
    Co = CN // NC
    y = yo.float().reshape(b, NC, T, Hp, Wp, 2, 2, Co).permute(0, 7, 1, 2, 3, 5, 4, 6)
    y = y.reshape(b, Co * NC, T, Hp * 2, Wp * 2)
    out.copy_(y[:, :, :Tx, :Hx, :Wx])
    return out


def swap01(src, dst, A, B, inner):
    dst.copy_(src.reshape(A, B, inner).transpose(0, 1).reshape(dst.shape))
    return dst
