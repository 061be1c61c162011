"""Shared test helpers (test infrastructure)."""
import os

import torch

from magicdrive_v2_b200 import synth
from magicdrive_v2_b200.config import MagicDriveSTDiT3Config
from magicdrive_v2_b200.stdit3 import MagicDriveSTDiT3

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel_l2(a, b):
    a, b = a.double(), b.double()
    return ((a - b).norm() / b.norm()).item()


def build_model(workload_name, seed=0, **extra_kw):
    wl = synth.WORKLOADS[workload_name]
    kw = synth.model_kwargs(**wl["model"])
    kw.update(extra_kw)
    torch.manual_seed(0)
    model = MagicDriveSTDiT3(MagicDriveSTDiT3Config(**kw)).eval()
    synth.fill_params(model, seed=seed)
    return model, wl, kw


def load_golden(name):
    return torch.load(os.path.join(GOLDEN, name), weights_only=False)


class use_ops:
    """Context manager that injects an ops backend into the drop-in (tests only)."""

    def __init__(self, backend):
        self.backend = backend

    def __enter__(self):
        self.prev = MagicDriveSTDiT3._ops
        MagicDriveSTDiT3._ops = self.backend

    def __exit__(self, *a):
        MagicDriveSTDiT3._ops = self.prev
