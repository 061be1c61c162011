"""Null-condition builders used by the samplers that run the conditional and unconditional halves of classifier-free
guidance as two separate denoiser calls.  Mirrors of ``magicdrivedit/utils/inference_utils.py:144-226``
(``replace_with_null_condition``) and ``:229-296`` (``add_null_condition``): same arguments, same result structure,
including the reference's quirk of dropping the last row of ``rel_pos`` (``[..., :-1, :]``) on the null side only."""
from __future__ import annotations

from copy import deepcopy

import torch

_UNCHANGED = ("mv_order_map", "t_order_map", "height", "width", "num_frames", "fps")


def _twice(t):                                         # einops "b ... -> (2 b) ..."
    return t.repeat(2, *([1] * (t.dim() - 1)))


def _filled(like, row):
    """``like``-shaped tensor whose trailing [3, n] matrices all equal ``row``."""
    return row.to(like.dtype)[None].expand(like.numel() // row.numel(), *row.shape).reshape(like.shape).clone()


def _null_of(key, value, uncond_cam, uncond_rel_pos):
    """-> (the conditional tensor as the reference pairs it, its null counterpart)."""
    if key == "bbox":
        return value, {k: torch.zeros_like(v) for k, v in value.items()}
    if key == "cams":
        return value, _filled(value, uncond_cam)                         # [B·NC, T, 1, 3, 7]
    if key == "rel_pos":
        value = value[..., :-1, :]                                       # [B·NC, T, 1, 4, 4] -> [..., 3, 4]
        return value, _filled(value, uncond_rel_pos)
    return value, torch.zeros_like(value)                                # maps


def _cat(a, b):
    if isinstance(a, dict):
        return {k: torch.cat([a[k], b[k]], dim=0) for k in a}
    return torch.cat([a, b], dim=0)


def replace_with_null_condition(_model_args, uncond_cam, uncond_rel_pos, uncond_y, keys, append=False):
    """New dict in which every condition named in ``keys`` is replaced by (``append=False``) or extended with
    (``append=True``) its null version; the other entries are deep-copied / repeated twice.  ``keys`` is consumed like
    in the reference; names left over raise ``RuntimeError``."""
    out, handled = {}, []
    if "y" in keys and "y" in _model_args:
        handled.append("y")
        out["y"] = torch.cat([_model_args["y"], uncond_y], 0) if append else uncond_y
        keys.remove("y")
    for key in ("bbox", "cams", "rel_pos", "maps"):
        if key in keys and key in _model_args:
            handled.append(key)
            cond, null = _null_of(key, _model_args[key], uncond_cam, uncond_rel_pos)
            out[key] = _cat(cond, null) if append else null
            keys.remove(key)
    if len(keys) > 0:
        raise RuntimeError(f"{keys} left unhandled with {_model_args.keys()}")
    for k, v in _model_args.items():
        if k in handled:
            continue
        if k in _UNCHANGED:
            out[k] = v
        elif k == "bbox":
            out[k] = {kb: _twice(vb) for kb, vb in v.items()}
        else:
            out[k] = _twice(v) if append else deepcopy(v)
    return out


def add_null_condition(_model_args, uncond_cam, uncond_rel_pos, prepend=False, use_map0=False):
    """Doubled conditions for a batch-doubled CFG call: boxes / cameras / frame poses (and maps with ``use_map0``)
    get their null version appended (or prepended), everything else is repeated twice."""
    out, handled = {}, []
    for key in ("bbox", "cams", "rel_pos") + (("maps",) if use_map0 else ()):
        if key in _model_args:
            handled.append(key)
            cond, null = _null_of(key, _model_args[key], uncond_cam, uncond_rel_pos)
            out[key] = _cat(null, cond) if prepend else _cat(cond, null)
    for k, v in _model_args.items():
        if k in handled:
            continue
        out[k] = v if k in _UNCHANGED else _twice(v)
    return out
