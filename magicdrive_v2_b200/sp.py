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


def ulysses_attention(ops, q, k, v, out, d, kv_map, scale, group, p):
    """q/k/v: views [N_local, C] of one qkv buffer (row stride 3C); out: [N_local, C]."""
    BT, Sl, H, D = d.B * d.T, d.S, d.H, d.D
    Hp = H // p
    qkv = torch.stack([q, k, v], dim=1)                                   # [N_l, 3, C] (copy)
    send = qkv.view(BT, Sl, 3, p, Hp * D).permute(3, 0, 1, 2, 4).contiguous()    # [p, BT, Sl, 3, HpD]
    recv = torch.empty_like(send)
    dist.all_to_all_single(recv, send, group=group)
    full = recv.permute(1, 0, 2, 3, 4).reshape(BT * p * Sl, 3, Hp * D)    # rows (i, r, sl) = (i, s)
    full = full.contiguous()
    o = torch.empty(BT * p * Sl, Hp * D, dtype=q.dtype, device=q.device)
    ops.attention(full[:, 0], full[:, 1], full[:, 2], o, BT, p * Sl, p * Sl, Hp, D, scale, kv_map=kv_map)
    send2 = o.view(BT, p, Sl, Hp * D).permute(1, 0, 2, 3).contiguous()     # [p(dst), BT, Sl, HpD]
    recv2 = torch.empty_like(send2)
    dist.all_to_all_single(recv2, send2, group=group)
    out.view(BT, Sl, p, Hp * D).copy_(recv2.permute(1, 2, 0, 3))           # heads back in rank order


def gather_tokens(yo, d, p, group):
    """[β·T·S/p, n] on every rank -> [β·T·S, n] (all-gather of the 64-channel final-layer output
    instead of the C-channel residual stream the reference gathers, magicdrive_stdit3.py:1202)."""
    n = yo.shape[1]
    parts = [torch.empty_like(yo) for _ in range(p)]
    dist.all_gather(parts, yo.contiguous(), group=group)
    full = torch.stack(parts, dim=0).view(p, d.B * d.T, d.S, n).permute(1, 0, 2, 3)
    return full.reshape(d.B * d.T * p * d.S, n).contiguous()
