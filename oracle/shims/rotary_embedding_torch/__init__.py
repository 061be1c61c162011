"""Shim restating rotary-embedding-torch==0.5.3 (requirements/requirements.txt:20), 'lang' freqs,
theta=1e4, no xpos, no interpolation. The package source is not on this box (no network): this
restatement IS the spec the oracle runs with; parity of RoPE is therefore pinned
to this shim, not to the upstream wheel."""
import torch
import torch.nn as nn
from einops import rearrange, repeat


class RotaryEmbedding(nn.Module):
    def __init__(self, dim, theta=10000):
        super().__init__()
        self.freqs = nn.Parameter(
            1.0 / (theta ** (torch.arange(0, dim, 2)[: dim // 2].float() / dim)), requires_grad=False)

    def rotate_queries_or_keys(self, t, seq_dim=-2, offset=0):
        # t: [..., T, D]
        pos = torch.arange(t.shape[seq_dim], device=t.device, dtype=t.dtype) + offset
        ang = torch.einsum("..., f -> ... f", pos.type(self.freqs.dtype), self.freqs)
        ang = repeat(ang, "... n -> ... (n r)", r=2)
        x1, x2 = rearrange(t, "... (d r) -> ... d r", r=2).unbind(-1)
        rot = rearrange(torch.stack((-x2, x1), -1), "... d r -> ... (d r)")
        return t * ang.cos() + rot * ang.sin()
