"""Small helpers the denoiser's callers and configs rely on: mirrors of
``magicdrivedit/models/magicdrive/utils.py:4-27`` (``ensure_kv_is_int``, ``zero_module``, ``load_module``),
``magicdrivedit/utils/misc.py:18-20`` (``warn_once``) and ``magicdrivedit/acceleration/checkpoint.py:19-24``
(``auto_grad_checkpoint``)."""
from __future__ import annotations

import importlib
import logging
from functools import lru_cache


def ensure_kv_is_int(view_pair: dict) -> dict:
    """yaml keys may be ints, json keys cannot: ``{"0": ["5", "1"]}`` -> ``{0: [5, 1]}`` (the ``mv_order_map``)."""
    return {int(k): [int(vi) for vi in v] for k, v in view_pair.items()}


def zero_module(module):
    """Zeroes every parameter of ``module`` in place and returns it."""
    for p in module.parameters():
        p.detach().zero_()
    return module


def load_module(name: str):
    """``"pkg.mod.Class"`` -> the class object.  Dotted names that point into the reference's embedder module resolve
    to the native classes of ``magicdrive_v2_b200.embedders`` (the configs name embedders this way,
    magicdrive_stdit3.py:531-543)."""
    p, m = name.rsplit(".", 1)
    if p.startswith("magicdrivedit."):
        from .embedders import resolve_embedder
        return resolve_embedder(name)
    return getattr(importlib.import_module(p), m)


@lru_cache(None)
def warn_once(msg: str):
    logging.warning(msg)


def auto_grad_checkpoint(module, *args, **kwargs):
    """Inference drop-in: activation checkpointing is a training feature, so this is the reference's call-through
    branch (``grad_checkpointing`` unset); a module flagged for checkpointing is refused rather than silently run."""
    if getattr(module, "grad_checkpointing", False):
        raise NotImplementedError("gradient checkpointing is outside the denoising hot path (DESIGN.md §7)")
    return module(*args, **kwargs)
