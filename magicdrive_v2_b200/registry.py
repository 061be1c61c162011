"""Registry boundary of the drop-in (reference: magicdrivedit/registry.py:7-46).

The reference builds the denoiser with ``build_module(cfg.model, MODELS, **extra)`` where MODELS is
an mmengine ``Registry``.  mmengine is not a dependency here: this is a dict-backed registry with
the same two calls (``register_module(name)`` decorator, ``build(cfg)``), so a config dict
``{"type": "MagicDriveSTDiT3-XL/2", ...}`` builds the B200-native module unchanged.
"""
from copy import deepcopy

import torch.nn as nn


class Registry:
    def __init__(self, name: str):
        self.name = name
        self._modules = {}

    def register_module(self, name=None, force=False, module=None):
        def _reg(obj):
            key = name if name is not None else obj.__name__
            if key in self._modules and not force:
                raise KeyError(f"{key} is already registered in {self.name}")
            self._modules[key] = obj
            return obj
        return _reg(module) if module is not None else _reg

    @staticmethod
    def _populate():
        """The reference's ``models/__init__`` fills the registry on import; here the model / sampler modules register
        themselves when imported, so every lookup imports them first (they load the CUDA library lazily: safe
        without libmdv2.so)."""
        from . import brushnet, rflow, stdit3  # noqa: F401

    def get(self, key):
        self._populate()
        return self._modules.get(key)

    def __getitem__(self, key):
        self._populate()
        if key not in self._modules:
            raise KeyError(f"{key} is not in the {self.name} registry")
        return self._modules[key]

    def __contains__(self, key):
        self._populate()
        return key in self._modules

    def build(self, cfg: dict, **kwargs):
        cfg = dict(cfg)
        cfg.update(kwargs)
        if "type" not in cfg:
            raise KeyError(f"`cfg` must contain the key \"type\", but got {cfg}")
        key = cfg.pop("type")
        self._populate()
        if key not in self._modules:
            raise KeyError(f"{key} is not in the {self.name} registry")
        return self._modules[key](**cfg)


def build_module(module, builder: Registry, **kwargs):
    """Same contract as registry.py:7-28: dict -> build, nn.Module -> itself, None -> None."""
    if module is None:
        return None
    if isinstance(module, dict):
        cfg = deepcopy(module)
        cfg.update(kwargs)
        return builder.build(cfg)
    if isinstance(module, nn.Module):
        return module
    raise TypeError(f"Only support dict and nn.Module, but got {type(module)}.")


MODELS = Registry("model")
SCHEDULERS = Registry("scheduler")
