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
