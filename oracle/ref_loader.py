"""TEST INFRASTRUCTURE — imports the REAL reference (``/root/reference``) behind the shims in
``oracle/shims`` so that golden vectors can be generated from the reference's own code.

In the build container it reads ``/root/reference``; on the GPU box the same unmodified files from ``oracle/_ref``
(``oracle/populate_ref.py``), where only ``bench.py``'s CPU legs use it.  Recipe: SURVEY.md App. C.
"""
import importlib
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
_SHIMS = os.path.join(_HERE, "shims")


def _find_root():
    """/root/reference in the build container; on the GPU box the git-ignored copy of the same unmodified files that
    oracle/populate_ref.py left under oracle/_ref (it ships with the snapshot like the built .so)."""
    env = os.environ.get("MDV2_REFERENCE_ROOT")
    for cand in ([env] if env else []) + ["/root/reference", os.path.join(_HERE, "_ref")]:
        if os.path.isdir(os.path.join(cand, "magicdrivedit")):
            return cand
    return env or "/root/reference"


REF_ROOT = _find_root()


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "magicdrivedit"))


def _ns(name, path):
    m = types.ModuleType(name)
    m.__path__ = [path]
    sys.modules[name] = m
    return m


def load_reference():
    """Returns the reference module ``magicdrivedit.models.magicdrive.magicdrive_stdit3``."""
    if not available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")
    name = "magicdrivedit.models.magicdrive.magicdrive_stdit3"
    if name in sys.modules:
        return sys.modules[name]
    os.environ["USE_XFORMERS"] = "False"
    if _SHIMS not in sys.path:
        sys.path.insert(0, _SHIMS)
    pkg = os.path.join(REF_ROOT, "magicdrivedit")
    # namespace stubs: skip the package __init__s that pull diffusers / ftfy / datasets
    _ns("magicdrivedit", pkg)
    _ns("magicdrivedit.models", os.path.join(pkg, "models"))
    _ns("magicdrivedit.models.vae", os.path.join(pkg, "models", "vae"))
    _ns("magicdrivedit.utils", os.path.join(pkg, "utils"))
    ck = types.ModuleType("magicdrivedit.utils.ckpt_utils")

    def load_checkpoint(*a, **k):
        raise NotImplementedError("oracle stub")

    ck.load_checkpoint = load_checkpoint
    sys.modules["magicdrivedit.utils.ckpt_utils"] = ck
    return importlib.import_module(name)


def build_reference_model(kwargs: dict, brushnet=False):
    """Builds the reference nn.Module from constructor kwargs (see synth.model_kwargs).  ``brushnet``: False | True |
    "sde" (the SDE-BrushNet class, magicdrive_stdit3.py:1870)."""
    ref = load_reference()
    cfg = ref.MagicDriveSTDiT3Config(**kwargs)
    cls = (ref.MagicDriveSTDiT3SDEBrushNet if brushnet == "sde" else
           ref.MagicDriveSTDiT3BrushNet if brushnet else ref.MagicDriveSTDiT3)
    return cls(cfg)


def load_reference_sampler():
    """Returns the reference module ``magicdrivedit.schedulers.rf`` (RFLOW, timestep_transform)."""
    load_reference()
    name = "magicdrivedit.schedulers.rf"
    if name in sys.modules:
        return sys.modules[name]
    pkg = os.path.join(REF_ROOT, "magicdrivedit")
    _ns("magicdrivedit.schedulers", os.path.join(pkg, "schedulers"))      # skip iddpm / dpm-solver imports
    # magicdrivedit/utils/inference_utils.py imports mmcv / datasets at module level; the samplers only need its two
    # null-condition builders, so exactly those two function definitions of the REAL file are compiled (from where the
    # file lies) into a stand-in module
    import ast
    path = os.path.join(pkg, "utils", "inference_utils.py")
    with open(path) as f:
        tree = ast.parse(f.read(), filename=path)
    wanted = ("replace_with_null_condition", "add_null_condition")
    tree.body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in wanted]
    assert len(tree.body) == len(wanted)
    iu = types.ModuleType("magicdrivedit.utils.inference_utils")
    import copy
    import torch
    from einops import repeat
    iu.__dict__.update(torch=torch, repeat=repeat, deepcopy=copy.deepcopy)
    exec(compile(tree, path, "exec"), iu.__dict__)
    sys.modules["magicdrivedit.utils.inference_utils"] = iu
    return importlib.import_module(name)
