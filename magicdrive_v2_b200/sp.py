"""Sequence parallelism for the STDiT3 hot path (reference: magicdrivedit/acceleration/
communications.py:11-66,118-155 and parallel_states.py; SP attention blocks.py:546-645).

Tokens stay sharded on S for the whole stack (rank r owns s in [r·S/p, (r+1)·S/p)).  Only the two
spatial attentions exchange data: one all-to-all turns the local ``[β·T, S/p, NH, D]`` q/k/v into
``[β·T, S, NH/p, D]`` (rank r owns heads [r·NH/p, …)), and one turns the attention output back.
q, k and v travel in ONE message (the reference sends q and kv separately and wraps each call in
split+cat copies); cross-attention needs no communication at all because K/V (the condition
tokens) are replicated — the reference's 164 all-to-alls there are dropped.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

_GROUPS = {}


def set_sequence_parallel_group(group):
    _GROUPS["sequence"] = group


def get_sequence_parallel_group():
    group = _GROUPS.get("sequence", None)
    if group is None:
        raise RuntimeError("sequence_parallel_group is None")
    return group


def set_data_parallel_group(group):
    _GROUPS["data"] = group


def get_data_parallel_group():
    group = _GROUPS.get("data", None)
    if group is None:
        raise RuntimeError("data_parallel_group is None")
    return group


def all_to_all(x: torch.Tensor, group, scatter_dim: int = 2, gather_dim: int = 1) -> torch.Tensor:
    """Same contract as communications.py:60-66 (forward only), on ``all_to_all_single``."""
    p = dist.get_world_size(group)
    if p == 1:
        return x
    assert x.shape[scatter_dim] % p == 0
    send = torch.stack(x.chunk(p, dim=scatter_dim)).contiguous()          # [p, ...]
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)
    return torch.cat(list(recv.unbind(0)), dim=gather_dim).contiguous()


def split(x: torch.Tensor, group, dim: int) -> torch.Tensor:
    """Local slice along ``dim`` (``_split``, communications.py:118-135)."""
    p, r = dist.get_world_size(group), dist.get_rank(group)
    if p == 1:
        return x
    assert x.size(dim) % p == 0
    return x.chunk(p, dim=dim)[r].contiguous()


def gather(x: torch.Tensor, group, dim: int) -> torch.Tensor:
    """all-gather + concat along ``dim`` (``_gather``, communications.py:138-155)."""
    p = dist.get_world_size(group)
    if p == 1:
        return x
    x = x.contiguous()
    parts = [torch.empty_like(x) for _ in range(p)]
    dist.all_gather(parts, x, group=group)
    return torch.cat(parts, dim=dim).contiguous()


_BUFS = {}


def _buf(tag, shape, like):
    """Per-(tag, shape) scratch reused across the 69 attention calls of a forward (stream-ordered reuse)."""
    key = (tag, tuple(shape), like.dtype, like.device)
    t = _BUFS.get(key)
    if t is None:
        if len(_BUFS) > 64:
            _BUFS.clear()
        t = _BUFS[key] = torch.empty(shape, dtype=like.dtype, device=like.device)
    return t


def ulysses_attention(ops, qkv, out, d, kv_map, scale, group, p):
    """Spatial / cross-view attention under sequence parallelism (blocks.py:546-645, without its copies).

    qkv: [N_l, 3C] local tokens (rows (β, t, s_l)), columns q|k|v with heads in rank order; out: [N_l, C].
    The work is split into the b scheduler-batch halves (each closed under the cross-view neighbour map), and
    the chunks are software-pipelined: while chunk c runs attention, the NCCL stream moves chunk c+1's q/k/v
    and chunk c-1's output.  Every re-ordering is one `swap01` launch at HBM speed:

        qkv [n·3][p][HpD]        -> send  [p][n·3][HpD]          heads scatter   (then all-to-all)
        recv [p][β'·T][Sl·3·HpD] -> full  [β'·T][p·Sl][3][HpD]   sequence gather
        o    [β'·T][p][Sl·HpD]   -> send2 [p][β'·T][Sl·HpD]      (then all-to-all)
        recv2 [p][n][HpD]        -> out   [n][p·HpD]             heads gather
    """
    Sl, H, D, C = d.S, d.H, d.D, d.C
    Hp = H // p
    HpD = Hp * D
    n_chunks = d.b if d.b > 1 else 1
    BTc = d.NC * d.T * (d.b // n_chunks)                 # sequences per chunk
    Nc = BTc * Sl                                        # local tokens per chunk
    assert qkv.is_contiguous() and out.is_contiguous() and qkv.shape == (d.N, 3 * C)
    maps = [None] * n_chunks
    if kv_map is not None:
        for c in range(n_chunks):
            m = kv_map[c * BTc:(c + 1) * BTc] - c * BTc
            maps[c] = m.contiguous()
    work1, work2 = [None] * n_chunks, [None] * n_chunks
    recv, recv2 = [None] * n_chunks, [None] * n_chunks
    for c in range(n_chunks):
        send = _buf(("send", c), (p, Nc * 3, HpD), qkv)
        recv[c] = _buf(("recv", c), (p, Nc * 3, HpD), qkv)
        ops.swap01(qkv[c * Nc:(c + 1) * Nc], send, Nc * 3, p, HpD)
        work1[c] = dist.all_to_all_single(recv[c], send, group=group, async_op=True)
    for c in range(n_chunks):
        full = _buf(("full", c), (BTc * p * Sl, 3 * HpD), qkv)
        o = _buf(("o", c), (BTc * p * Sl, HpD), qkv)
        send2 = _buf(("send2", c), (p, Nc, HpD), qkv)
        recv2[c] = _buf(("recv2", c), (p, Nc, HpD), qkv)
        work1[c].wait()
        ops.swap01(recv[c], full, p, BTc, Sl * 3 * HpD)
        ops.attention(full[:, :HpD], full[:, HpD:2 * HpD], full[:, 2 * HpD:], o, BTc, p * Sl, p * Sl, Hp, D, scale,
                      kv_map=maps[c])
        ops.swap01(o, send2, BTc, p, Sl * HpD)
        work2[c] = dist.all_to_all_single(recv2[c], send2, group=group, async_op=True)
    for c in range(n_chunks):
        work2[c].wait()
        ops.swap01(recv2[c], out[c * Nc:(c + 1) * Nc], p, Nc, HpD)


def gather_tokens(yo, d, p, group):
    """[β·T·S/p, n] on every rank -> [β·T·S, n] (all-gather of the 64-channel final-layer output
    instead of the C-channel residual stream the reference gathers, magicdrive_stdit3.py:1202)."""
    n = yo.shape[1]
    parts = [torch.empty_like(yo) for _ in range(p)]
    dist.all_gather(parts, yo.contiguous(), group=group)
    full = torch.stack(parts, dim=0).view(p, d.B * d.T, d.S, n).permute(1, 0, 2, 3)
    return full.reshape(d.B * d.T * p * d.S, n).contiguous()
