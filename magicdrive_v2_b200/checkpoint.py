"""Checkpoint loading with the reference's accepted formats (``magicdrivedit/utils/ckpt_utils.py:122-143``):
a ``.pt/.pth`` state dict (with the ``reparameter`` fix-ups of ``:33-64``), a ``.safetensors`` file, or a
ColossalAI ``GeneralCheckpointIO`` sharded directory (``<dir>/<model_name>/`` holding an index json plus
``*.bin`` / ``*.safetensors`` shards, ``:80-82``).  Only loading is in scope (inference path)."""
from __future__ import annotations

import json
import logging
import os

import torch

logger = logging.getLogger(__name__)


def reparameter(ckpt: dict, model) -> dict:
    """Drops stale position tables and resizes ``y_embedder.y_embedding`` to the model's text length
    by repeating / cutting the last row (ckpt_utils.py:33-64)."""
    ckpt.pop("pos_embed_temporal", None)
    ckpt.pop("pos_embed", None)
    key = "y_embedder.y_embedding"
    if key in ckpt:
        have, want = ckpt[key].shape[0], model.y_embedder.y_embedding.shape[0]
        if have < want:
            logger.info("Extend y_embedding from %s to %s", have, want)
            extra = ckpt[key][-1:].expand(want - have, -1).clone()
            ckpt[key] = torch.cat([ckpt[key], extra], dim=0)
        elif have > want:
            logger.info("Shrink y_embedding from %s to %s", have, want)
            ckpt[key] = ckpt[key][:want]
    return ckpt


def _load_file(path: str) -> dict:
    if path.endswith(".safetensors"):
        from safetensors.torch import load_file
        return load_file(path)
    return torch.load(path, map_location="cpu", weights_only=True)


def _load_sharded_dir(model, ckpt_dir: str, strict: bool):
    """ColossalAI GeneralCheckpointIO layout: an ``*.index.json`` with a ``weight_map`` (name -> shard file),
    or a single un-sharded weight file."""
    index = [f for f in sorted(os.listdir(ckpt_dir)) if f.endswith(".index.json")]
    if index:
        with open(os.path.join(ckpt_dir, index[0])) as f:
            weight_map = json.load(f)["weight_map"]
        state = {}
        for shard in sorted(set(weight_map.values())):
            state.update(_load_file(os.path.join(ckpt_dir, shard)))
    else:
        files = [f for f in sorted(os.listdir(ckpt_dir)) if f.endswith((".safetensors", ".bin", ".pt"))]
        if not files:
            raise ValueError(f"no weight files under {ckpt_dir}")
        state = {}
        for f in files:
            state.update(_load_file(os.path.join(ckpt_dir, f)))
    return model.load_state_dict(state, strict=strict)


def load_checkpoint(model, ckpt_path: str, model_name: str = "model", strict: bool = False):
    if ckpt_path.endswith((".pt", ".pth")):
        assert os.path.isfile(ckpt_path), f"Could not find DiT checkpoint at {ckpt_path}"
        state = reparameter(_load_file(ckpt_path), model)
        res = model.load_state_dict(state, strict=strict)
    elif ckpt_path.endswith(".safetensors"):
        res = model.load_state_dict(_load_file(ckpt_path), strict=False)      # the reference is lenient here
    elif os.path.isdir(ckpt_path):
        res = _load_sharded_dir(model, os.path.join(ckpt_path, model_name), strict)
    else:
        raise ValueError(f"Invalid checkpoint path: {ckpt_path}")
    logger.info("Missing keys: %s", res.missing_keys)
    logger.info("Unexpected keys: %s", res.unexpected_keys)
    return res


def _as_state(src) -> dict:
    if isinstance(src, dict):
        return src
    if os.path.isdir(src):
        files = [f for f in sorted(os.listdir(src)) if f.endswith((".safetensors", ".bin", ".pt", ".pth"))]
        if not files:
            raise ValueError(f"no weight files under {src}")
        state = {}
        for f in files:
            state.update(_load_file(os.path.join(src, f)))
        return state
    return _load_file(src)


def _sub(state: dict, prefix: str) -> dict:
    return {k[len(prefix):]: v for k, v in state.items() if k.startswith(prefix)}


def _init_from_base(model, state: dict, spatial: str, temporal: str | None, what: str):
    """Shared body of the two initialisers below: helper modules by name (non-strict), ``x_control_embedder`` from the
    base ``x_embedder``, then the base model's block stack into BOTH the base and the control streams."""
    res = model.load_state_dict(state, strict=False)
    if getattr(model, "x_control_embedder", None) is not None:
        model.x_control_embedder.load_state_dict(_sub(state, "x_embedder."))
    skip_m, skip_u = ("base_blocks_", "control_blocks_"), tuple(p for p in (spatial, temporal) if p)
    logger.info("1st, Load from %s with missing=%s, unexpected=%s", what,
                [k for k in res.missing_keys if not k.startswith(skip_m)],
                [k for k in res.unexpected_keys if not k.startswith(skip_u)])
    sp_state = _sub(state, spatial + ".")
    report = {}
    report["base_s"] = model.base_blocks_s.load_state_dict(sp_state, strict=False)
    n_ctrl = len(model.control_blocks_s)
    ctrl_state = {k: v for k, v in sp_state.items() if int(k.split(".", 1)[0]) < n_ctrl}
    report["control_s"] = model.control_blocks_s.load_state_dict(ctrl_state, strict=False)
    if temporal is not None:
        t_state = _sub(state, temporal + ".")
        if model.base_blocks_t is not None:
            report["base_t"] = model.base_blocks_t.load_state_dict(t_state, strict=False)
        if model.control_blocks_t is not None:
            ct_state = {k: v for k, v in t_state.items() if int(k.split(".", 1)[0]) < len(model.control_blocks_t)}
            report["control_t"] = model.control_blocks_t.load_state_dict(ct_state, strict=False)
    for name, r in report.items():
        logger.info("Load %s from %s with missing=%s, unexpected=%s", name, what, r.missing_keys, r.unexpected_keys)
    return model


def load_from_stdit3_pretrained(model, from_pretrained):
    """``magicdrive_stdit3.py:2473-2514`` at state-dict level: initialise from an OpenSora STDiT3 checkpoint (a local
    file / directory or an already loaded state dict) — ``spatial_blocks.*`` / ``temporal_blocks.*`` go into both the
    base and the control streams, the remaining modules load by name.  (The reference instantiates the base
    ``STDiT3`` class to obtain its state dict; only the weights matter.)"""
    return _init_from_base(model, _as_state(from_pretrained), "spatial_blocks", "temporal_blocks",
                           from_pretrained if isinstance(from_pretrained, str) else "<state dict>")


def load_from_pixart_pretrained(model, pretrained):
    """``magicdrive_stdit3.py:2517-2552`` at state-dict level: initialise the spatial streams from a PixArt-XL/2
    checkpoint (``blocks.*``; the control stream takes the first ``control_depth`` blocks)."""
    return _init_from_base(model, _as_state(pretrained), "blocks", None, pretrained if isinstance(pretrained, str) else "<state dict>")
