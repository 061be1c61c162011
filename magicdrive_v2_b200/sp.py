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

import logging

import torch
import torch.distributed as dist

logger = logging.getLogger(__name__)

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


def ulysses_attention(ops, qkv, out, d, kv_map, scale, group, p, produce, consume):
    """Spatial / cross-view attention under sequence parallelism (blocks.py:546-645, without its copies).

    qkv: [N_l, 3C] local tokens (rows (β, t, s_l)), columns q|k|v with heads in rank order; out: [N_l, C].
    ``produce(rows, tab)`` computes ``qkv[rows]`` and ``consume(rows, tab)`` folds ``out[rows]`` into the
    residual stream (the projections around the attention, see MagicDriveSTDiT3._spatial_attention).

    The work is split into the b scheduler-batch halves (each closed under the cross-view neighbour map) and
    software-pipelined over them, with NCCL running on its own stream:

        produce(0) | produce(1) ‖ a2a(0) | attn(0) ‖ a2a(1) | attn(1) ‖ a2a'(0) | consume(0) ‖ a2a'(1) | consume(1)

    so every all-to-all has a GEMM or an attention of the other half to hide behind.  Every re-ordering is one
    `swap01` launch at HBM speed:

        qkv [n·3][p][HpD]        -> send  [p][n·3][HpD]          heads scatter   (then all-to-all)
        recv [p][β'·T][Sl·3·HpD] -> full  [β'·T][p·Sl][3][HpD]   sequence gather
        o    [β'·T][p][Sl·HpD]   -> send2 [p][β'·T][Sl·HpD]      (then all-to-all)
        recv2 [p][n][HpD]        -> out   [n][p·HpD]             heads gather
    """
    _log_path("NCCL all_to_all_single with swap01 pack / unpack")
    Sl, H, D, C = d.S, d.H, d.D, d.C
    Hp = H // p
    HpD = Hp * D
    n_chunks = d.b if d.b > 1 else 1
    bc = d.b // n_chunks                                 # scheduler-batch entries per chunk
    BTc = d.NC * d.T * bc                                # sequences per chunk
    Nc = BTc * Sl                                        # local tokens per chunk
    assert qkv.is_contiguous() and out.is_contiguous() and qkv.shape == (d.N, 3 * C)
    maps = [None] * n_chunks
    if kv_map is not None:
        for c in range(n_chunks):
            m = kv_map[c * BTc:(c + 1) * BTc] - c * BTc
            maps[c] = m.contiguous()

    def rows_of(c):
        return slice(c * Nc, (c + 1) * Nc)

    def tab_of(c):                                       # modulation tables are [b·T, K, C]
        return lambda m: m[c * bc * d.T:(c + 1) * bc * d.T]

    work1, work2 = [None] * n_chunks, [None] * n_chunks
    recv, recv2 = [None] * n_chunks, [None] * n_chunks
    for c in range(n_chunks):
        produce(rows_of(c), tab_of(c))
        send = _buf(("send", c), (p, Nc * 3, HpD), qkv)
        recv[c] = _buf(("recv", c), (p, Nc * 3, HpD), qkv)
        ops.swap01(qkv[rows_of(c)], send, Nc * 3, p, HpD)
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
        ops.swap01(recv2[c], out[rows_of(c)], p, Nc, HpD)
        consume(rows_of(c), tab_of(c))


# ---------------------------------------------------------------------------------------------------
# NVLink peer-memory exchange (one process per GPU of one NVSwitch box): the re-sharding all-to-alls become
# ONE libmdv2 scatter kernel each, storing straight into the final layout in the peers' HBM.
# ---------------------------------------------------------------------------------------------------
_SYMM = {}
_SYMM_OK = None


def peer_exchange_available(like: torch.Tensor, group) -> bool:
    """True when the ranks of ``group`` can map each other's device memory (torch symmetric memory over CUDA
    IPC / NVLink) and there are at most 8 of them.  ``MDV2_SP_NCCL=1`` forces the NCCL all-to-all path."""
    global _SYMM_OK
    import os
    if not like.is_cuda or os.environ.get("MDV2_SP_NCCL") or dist.get_world_size(group) > 8:
        return False
    if _SYMM_OK is None:
        try:
            import torch.distributed._symmetric_memory as symm
            t = symm.empty(1024, dtype=torch.uint8, device=like.device)
            symm.rendezvous(t, group.group_name)
            _SYMM_OK = True
        except Exception as e:                            # noqa: BLE001 - any failure means "not available here"
            _SYMM_OK = False
            logger.warning("symmetric memory unavailable (%s: %s): sequence parallelism falls back to the NCCL "
                           "all-to-all path", type(e).__name__, e)
        flag = torch.tensor([1 if _SYMM_OK else 0], device=like.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)      # all ranks take the same path
        _SYMM_OK = bool(flag.item())
    return _SYMM_OK


def symm_buffer(tag, shape, like, group):
    """-> (tensor, handle, peer base addresses): a symmetric allocation per (tag, shape), made collectively."""
    import torch.distributed._symmetric_memory as symm
    key = (tag, tuple(shape), like.dtype, like.device, group.group_name)
    ent = _SYMM.get(key)
    if ent is None:
        t = symm.empty(tuple(shape), dtype=like.dtype, device=like.device)
        hdl = symm.rendezvous(t, group.group_name)
        ent = _SYMM[key] = (t, hdl, [int(a) for a in hdl.buffer_ptrs])
    return ent


_PATH_LOGGED = set()


def _log_path(what):
    """One line per process and path: which exchange the sequence-parallel attention takes."""
    if what not in _PATH_LOGGED:
        _PATH_LOGGED.add(what)
        logger.info("sequence-parallel exchange: %s", what)


def exchange_path() -> str:
    """The exchange path(s) taken so far in this process (bench.py prints it in ``config``)."""
    return "; ".join(sorted(_PATH_LOGGED)) or "none yet"


stats = {"fused_qkv_scatter": 0, "fused_out_scatter": 0, "scatter_kernel": 0}     # which exchange ran (tests check the path taken)


def _fused_scatter(ops, qkv, D, C, p) -> bool:
    """GEMM-epilogue scatter of q|k|v (``MDV2_SP_FUSED=0`` keeps the separate scatter kernel)."""
    import os
    if os.environ.get("MDV2_SP_FUSED", "1") == "0" or not hasattr(ops, "gemm_qkv_scatter"):
        return False
    return (qkv.dtype == torch.bfloat16 and D in (64, 72) and (C // p) % D == 0 and C % 64 == 0
            and getattr(ops, "default_backend", 0) != getattr(ops, "BACKEND_SIMT", 1))


def ulysses_attention_peer(ops, qkv, ao_sym, d, kv_map, scale, group, p, rank, produce, consume):
    """Same contract as ``ulysses_attention`` over peer memory.  ``ao_sym = (tensor, handle, peer addresses)`` is
    the symmetric buffer that ``consume`` reads the attention output from (the workspace's ``ao``).

        produce (+ scatter of q|k|v heads -> owner ranks, straight into [β·T, S, 3, Hp·D], fused into the qkv
                 GEMM's epilogue when the tcgen05 kernel takes the shape) -> barrier -> attention
                -> scatter(output rows -> owner ranks, straight into their [N_l, C]) -> barrier -> consume

    No pack / unpack copies and no NCCL call: the scatter kernel reads local HBM once and writes the peers'
    HBM through NVLink.  The only cross-rank ordering is the barrier behind each scatter; a buffer is next
    overwritten only after a later barrier that every rank reaches after it has finished reading it."""
    Sl, H, D, C = d.S, d.H, d.D, d.C
    Hp = H // p
    HpD = Hp * D
    BT = d.B * d.T
    rows = slice(0, d.N)
    ident = lambda m: m                                   # noqa: E731
    full, hdl_f, ptrs_f = symm_buffer("full", (BT * p * Sl, 3 * HpD), qkv, group)
    fused = _fused_scatter(ops, qkv, D, C, p) and getattr(produce, "fusable", True)
    _log_path("fused GEMM / attention epilogue scatter over NVLink peer memory" if fused else
              "scatter kernels over NVLink peer memory")
    if fused:
        # the qkv projection's epilogue stores every head group straight into its owner's [β·T, S, 3, Hp·D] buffer:
        # GEMM and all-to-all are one kernel, the NVLink writes ride under the MMAs
        produce(rows, ident, scatter=(ptrs_f, Sl, rank))
        stats["fused_qkv_scatter"] += 1
    else:
        produce(rows, ident)
        stats["scatter_kernel"] += 1
        # index order (β·t, s, rank, q|k|v): the three runs of one (token, rank) land back to back in the peer's row,
        # so consecutive threads write 3·Hp·D contiguous elements over NVLink
        ops.scatter_runs(qkv, ptrs_f, 2, (BT, Sl, p, 3), (Sl * 3 * C, 3 * C, HpD, C),
                         (p * Sl * 3 * HpD, 3 * HpD, 0, HpD), HpD, dst_base=rank * Sl * 3 * HpD)
    hdl_f.barrier()
    o = _buf(("o", 0), (BT * p * Sl, HpD), qkv)
    ao, hdl_a, ptrs_a = ao_sym
    if fused and hasattr(ops, "attention_scatter") and ops.attention_scatter_supported(qkv, D, p):
        # the attention epilogue stores every output row into its owner's [N_l, C] buffer (o: partial sums of the
        # cross-view K/V sets only)
        ops.attention_scatter(full[:, :HpD], full[:, HpD:2 * HpD], full[:, 2 * HpD:], o, BT, p * Sl, p * Sl, Hp, D,
                              scale, ptrs_a, Sl, C, rank * HpD, kv_map=kv_map)
        stats["fused_out_scatter"] += 1
    else:
        ops.attention(full[:, :HpD], full[:, HpD:2 * HpD], full[:, 2 * HpD:], o, BT, p * Sl, p * Sl, Hp, D, scale,
                      kv_map=kv_map)
        ops.scatter_runs(o, ptrs_a, 1, (BT, p, Sl, 1), (p * Sl * HpD, Sl * HpD, HpD, 0), (Sl * C, 0, C, 0), HpD,
                         dst_base=rank * HpD)
    hdl_a.barrier()
    consume(rows, ident)


def gather_tokens(yo, d, p, group):
    """[β·T·S/p, n] on every rank -> [β·T·S, n] (all-gather of the 64-channel final-layer output
    instead of the C-channel residual stream the reference gathers, magicdrive_stdit3.py:1202)."""
    n = yo.shape[1]
    parts = [torch.empty_like(yo) for _ in range(p)]
    dist.all_gather(parts, yo.contiguous(), group=group)
    full = torch.stack(parts, dim=0).view(p, d.B * d.T, d.S, n).permute(1, 0, 2, 3)
    return full.reshape(d.B * d.T * p * d.S, n).contiguous()
