"""mdv2_attention_scatter_static — the no-allocation entry point (scratch in a static device array) — against the
existing mdv2_attention_scatter, through the C ABI.  Collected last on purpose."""
import ctypes
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _ops():
    from magicdrive_v2_b200 import ops
    return ops


def _rnd(*shape, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(*shape, generator=g).to(torch.bfloat16).to(DEV)


def _case(p, Hp, D, BT, Sl, n_sets, seed):
    HpD, C, S = Hp * D, p * Hp * D, p * Sl
    full = _rnd(BT * S, 3 * HpD, seed=seed)
    kv_map = None
    if n_sets == 2:
        kv_map = torch.tensor([[(i + 1) % BT, (i - 1) % BT] for i in range(BT)], dtype=torch.int32, device=DEV)
    return full[:, :HpD], full[:, HpD:2 * HpD], full[:, 2 * HpD:], kv_map, HpD, C, S


@pytest.mark.parametrize("p,Hp,D,BT,Sl,n_sets", [(2, 2, 72, 4, 333, 2), (4, 2, 64, 2, 100, 1), (1, 4, 72, 3, 200, 2)])
def test_static_scratch_matches_caller_scratch_two_calls_in_a_row(p, Hp, D, BT, Sl, n_sets):
    """Bit-identical to mdv2_attention_scatter (same kernel, same arithmetic) on two consecutive calls with
    different inputs: the second call finds the first call's partial sums in the shared array."""
    ops = _ops()
    rank = p - 1
    for call in range(2):
        q, k, v, kv_map, HpD, C, S = _case(p, Hp, D, BT, Sl, n_sets, seed=41 + call)
        want = [torch.zeros(BT * Sl, C, dtype=torch.bfloat16, device=DEV) for _ in range(p)]
        part = torch.empty(BT * S, HpD, dtype=torch.bfloat16, device=DEV)
        ops.attention_scatter(q, k, v, part, BT, S, S, Hp, D, D ** -0.5, [t.data_ptr() for t in want], Sl, C,
                              rank * HpD, kv_map=kv_map)
        got = [torch.zeros(BT * Sl, C, dtype=torch.bfloat16, device=DEV) for _ in range(p)]
        ptrs = [t.data_ptr() for t in got]
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        ops.attention_scatter_static(q, k, v, BT, S, S, Hp, D, D ** -0.5, ptrs, Sl, C, rank * HpD, kv_map=kv_map)
        assert torch.cuda.memory_allocated() == before          # nothing went through torch's allocator either
        torch.cuda.synchronize()
        for g in range(p):
            assert torch.equal(got[g], want[g]), (call, g)
        assert got[0][:, rank * HpD:(rank + 1) * HpD].abs().sum() > 0


def test_static_scratch_rejects_sizes_beyond_the_maximum():
    ops = _ops()
    lim = ops.static_scratch_bytes()
    assert lim == 16 * 1024 * 1024
    H, D, Lq = 16, 72, 5600
    n_seq = lim // (Lq * H * D * 2) + 1                         # one sequence too many
    q = torch.zeros(8, H * D, dtype=torch.bfloat16, device=DEV)  # never read: the size check comes first
    out = torch.zeros(8, H * D, dtype=torch.bfloat16, device=DEV)
    kv_map = torch.zeros(n_seq, 2, dtype=torch.int32, device=DEV)
    with pytest.raises(RuntimeError, match="exceed the static scratch"):
        ops.attention_scatter_static(q, q, q, n_seq, Lq, Lq, H, D, D ** -0.5, [out.data_ptr()], Lq, H * D, 0,
                                     kv_map=kv_map)


def test_static_entry_point_is_exported_by_the_c_abi():
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = ctypes.CDLL(os.path.join(root, "magicdrive_v2_b200", "libmdv2.so"))
    assert hasattr(lib, "mdv2_attention_scatter_static") and hasattr(lib, "mdv2_static_scratch_bytes")
