"""Integer index / permutation maps of the STDiT3 hot path.

These are the only parts of the reference's many ``rearrange`` copies that survive in the native
layout ``[β, T, S, C]``; they must be bit-exact and are tested against the oracle as such.
"""
from __future__ import annotations

import torch


def dynamic_size(T, H, W, patch=(1, 2, 2)):
    """ceil-div by the patch (magicdrive_stdit3.py:838-849)."""
    return tuple(-(-d // p) for d, p in zip((T, H, W), patch))


def h_pad_for(H0: int, W0: int, q) -> int:
    """P3: extra patch rows so that S % q == 0 (magicdrive_stdit3.py:1061-1063 / :1090-1093)."""
    if q is None or q <= 1 or (H0 * W0) % q == 0:
        return 0
    return q - H0 % q


def cross_view_kv_map(mv_order_map: dict, b: int, T: int, NC: int) -> torch.Tensor:
    """P5: for every (β, t) query sequence the two neighbour sequences whose K/V it attends to.

    Query sequence index i = β·T + t with β = bi·NC + v.  Pair j of target view v uses the K/V rows
    of view ``mv_order_map[v][j]`` of the same (bi, t) (magicdrive_stdit3.py:175-187, :304-322).
    Returns int32 [β·T, n_nbr] of sequence indices."""
    n_nbr = {len(v) for v in mv_order_map.values()}
    assert len(n_nbr) == 1, "every view needs the same number of neighbours"
    n_nbr = n_nbr.pop()
    assert sorted(mv_order_map.keys()) == list(range(NC))
    out = torch.empty(b * NC * T, n_nbr, dtype=torch.int32)
    for bi in range(b):
        for v in range(NC):
            for t in range(T):
                i = (bi * NC + v) * T + t
                for j, nb in enumerate(mv_order_map[v]):
                    out[i, j] = (bi * NC + int(nb)) * T + t
    return out


def sp_slice(S: int, rank: int, p: int):
    """P6: rank r owns s in [r·S/p, (r+1)·S/p) (communications.py:132-133)."""
    assert S % p == 0, f"S={S} should be divisible by {p}!"
    return rank * (S // p), (rank + 1) * (S // p)


def sp_head_slice(NH: int, rank: int, p: int):
    """P6: inside the spatial attentions rank r owns heads [r·NH/p, (r+1)·NH/p)."""
    assert NH % p == 0
    return rank * (NH // p), (rank + 1) * (NH // p)
