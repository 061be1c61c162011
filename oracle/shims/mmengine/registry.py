"""Shim of mmengine.registry.Registry: dict-backed register_module / build."""


class Registry:
    def __init__(self, name, locations=None, **kwargs):
        self.name = name
        self._module_dict = {}

    def register_module(self, name=None, force=False, module=None):
        def _register(obj):
            key = name if name is not None else obj.__name__
            self._module_dict[key] = obj
            return obj
        if module is not None:
            return _register(module)
        return _register

    def get(self, key):
        return self._module_dict.get(key)

    def build(self, cfg, **kwargs):
        cfg = dict(cfg)
        cfg.update(kwargs)
        obj = self._module_dict[cfg.pop("type")]
        return obj(**cfg)
