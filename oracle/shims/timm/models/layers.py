"""Shim of timm.models.layers.DropPath (identity at drop_prob=0 / eval)."""
import torch.nn as nn


class DropPath(nn.Module):
    def __init__(self, drop_prob: float = 0.0, scale_by_keep: bool = True):
        super().__init__()
        self.drop_prob = drop_prob

    def forward(self, x):
        if self.drop_prob == 0.0 or not self.training:
            return x
        raise NotImplementedError("oracle shim: stochastic depth is training-only")
