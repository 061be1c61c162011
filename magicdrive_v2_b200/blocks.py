"""Parameter holders + kernel schedule of one ``MultiViewSTDiT3Block``
(reference: magicdrive_stdit3.py:47-358; layers: blocks.py:383-514, :645-747).

The module keeps the reference's parameter names (``attn.qkv.weight`` …) so checkpoints load with
``strict=True``; its ``run`` method is the native schedule: every step is one call into the sm_100a
library (``ops``) operating in place on the canonical ``[β, T, S, C]`` residual stream.  There is no
eager/PyTorch path here.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn

from .embedders import CrossAttnParams, MlpParams, SelfAttnParams


@dataclass
class Dims:
    """Token-grid geometry of the (local) residual stream."""
    b: int          # scheduler batch (2 with CFG)
    NC: int         # views
    T: int          # latent frames
    S: int          # patch tokens per frame held by this rank
    C: int
    H: int          # heads
    D: int          # head dim

    @property
    def B(self):    # model batch β
        return self.b * self.NC

    @property
    def N(self):    # local tokens
        return self.B * self.T * self.S


def W(lin):
    """Effective weight of a Linear holder: the LoRA-folded copy when the BrushNet variant attached one
    (brushnet.py, magicdrive_stdit3.py:1422-1464), else the parameter itself."""
    w = getattr(lin, "_w_eff", None)
    return lin.weight if w is None else w


class Workspace:
    """Scratch activations shared by all blocks of one forward (allocated once per shape)."""

    def __init__(self, dims: Dims, dtype, device, mlp_ratio: float, n_sp: int = 1):
        N, C = dims.N, dims.C
        self.xm = torch.empty(N, C, dtype=dtype, device=device)
        self.qkv = torch.empty(N, 3 * C, dtype=dtype, device=device)
        self.ao = torch.empty(N, C, dtype=dtype, device=device)
        self.hid = torch.empty(N, int(C * mlp_ratio), dtype=dtype, device=device)
        self.key = (N, C, dtype, device, mlp_ratio)


class STDiT3Block(nn.Module):
    def __init__(self, hidden_size, num_heads, mlp_ratio=4.0, qk_norm=True, temporal=False,
                 is_control_block=False, skip_cross_attn=False, skip_cross_view=False):
        super().__init__()
        C = hidden_size
        self.hidden_size, self.num_heads = C, num_heads
        self.temporal = temporal
        self.is_control_block = is_control_block
        self.skip_cross_attn = skip_cross_attn
        self.skip_cross_view = skip_cross_view or temporal
        self.attn = SelfAttnParams(C, num_heads, qkv_bias=True, qk_norm=qk_norm)
        if not skip_cross_attn:
            self.cross_attn = CrossAttnParams(C, num_heads)
        self.mlp = MlpParams(C, int(C * mlp_ratio))
        if not self.skip_cross_view:
            self.cross_view_attn = SelfAttnParams(C, num_heads, qkv_bias=False, qk_norm=True)
            self.mva_proj = nn.Linear(C, C)
            nn.init.zeros_(self.mva_proj.weight)
            nn.init.zeros_(self.mva_proj.bias)
        else:
            self.mva_proj = None
        self.scale_shift_table = nn.Parameter(torch.randn(6, C) / C ** 0.5)
        if not self.skip_cross_view:
            self.scale_shift_table_mva = nn.Parameter(torch.randn(3, C) / C ** 0.5)
        if is_control_block:
            self.after_proj = nn.Linear(C, C)
            nn.init.zeros_(self.after_proj.weight)
            nn.init.zeros_(self.after_proj.bias)
        else:
            self.after_proj = None

    # ------------------------------------------------------------------------------------------
    # native schedule
    # ------------------------------------------------------------------------------------------
    def run(self, ops, x, ws: Workspace, d: Dims, mod, mod_mva, y_kv, ctx, skip_into=None):
        """x: [N, C] residual stream (updated in place).

        mod      fp32 [b·T, 6, C]   shift/scale/gate (msa), shift/scale/gate (mlp), x_mask-selected
        mod_mva  fp32 [b·T, 3, C]   cross-view modulation (base spatial blocks)
        y_kv     [β·Tc·L, 2C]       this block's kv_linear(y) (timestep-independent, cached)
        ctx      forward-wide constants: kv_map, rope tables, L, sequence-parallel hooks
        skip_into  residual stream that receives ``after_proj(x)`` (control blocks)
        """
        C, H, D, T, S, NC = d.C, d.H, d.D, d.T, d.S, d.NC
        geo = (T, S, NC)
        a = self.attn
        scale = D ** -0.5
        qn, kn = getattr(a.q_norm, "weight", None), getattr(a.k_norm, "weight", None)

        # (b) LN -> modulate ; (c) self-attention ; (d) gated residual.  The projections on either side of a
        # spatial attention are handed over as row-range callbacks (see MagicDriveSTDiT3._spatial_attention).
        def qkv_rows(rows, tab, scatter=None):
            ops.ln_modulate(x[rows], ws.xm[rows], tab(mod), 0, 1, geo)
            if scatter is not None:                     # sequence parallel: the epilogue stores into the peers' HBM
                ops.gemm_qkv_scatter(ws.xm[rows], W(a.qkv), a.qkv.bias, (qn, kn, D, C, C), *scatter)
            else:
                ops.gemm(ws.xm[rows], W(a.qkv), a.qkv.bias, ws.qkv[rows], geo=geo, rms=(qn, kn, D, C, C))

        # the fused projection + NVLink scatter needs the q / k RMSNorm epilogue (mdv2_gemm_qkv_scatter): with
        # qk_norm=False the sequence-parallel path keeps the plain GEMM and the separate scatter kernel
        qkv_rows.fusable = qn is not None and kn is not None

        def proj_rows(rows, tab):
            ops.gemm(ws.ao[rows], W(a.proj), a.proj.bias, x[rows], geo=geo, gate=(tab(mod), 2), res=x[rows])

        if self.temporal:
            rows = slice(0, x.shape[0])
            qkv_rows(rows, lambda m: m)
            ops.temporal_attention(ws.qkv, ctx.rope_cos, ctx.rope_sin, ws.ao, d.B, T, S, H, D, scale)
            proj_rows(rows, lambda m: m)
        else:
            ctx.spatial_attention(ops, ws.qkv, ws.ao, d, None, scale, qkv_rows, proj_rows)

        # (e) cross-attention onto the condition tokens: un-normed, un-gated
        if not self.skip_cross_attn:
            ca = self.cross_attn
            ops.gemm(x, W(ca.q_linear), ca.q_linear.bias, ws.xm, geo=geo)
            yk, yv = y_kv[:, :C], y_kv[:, C:]
            if ctx.y_frames == 1:
                ops.attention(ws.xm, yk, yv, ws.ao, d.B, T * S, ctx.L, H, D, scale)
            else:
                ops.attention(ws.xm, yk, yv, ws.ao, d.B * T, S, ctx.L, H, D, scale)
            ops.gemm(ws.ao, W(ca.proj), ca.proj.bias, x, geo=geo, res=x)

        # (f) cross-view attention (spatial base blocks): each view attends to its neighbours
        if not self.skip_cross_view:
            cv = self.cross_view_attn
            n_nbr = float(ctx.kv_map.shape[1])

            def cv_qkv_rows(rows, tab, scatter=None):
                ops.ln_modulate(x[rows], ws.xm[rows], tab(mod_mva), 0, 1, geo)
                rms = (cv.q_norm.weight, cv.k_norm.weight, D, C, C)
                if scatter is not None:
                    ops.gemm_qkv_scatter(ws.xm[rows], cv.qkv.weight, None, rms, *scatter)
                else:
                    ops.gemm(ws.xm[rows], cv.qkv.weight, None, ws.qkv[rows], geo=geo, rms=rms)

            def cv_proj_rows(rows, tab):
                # proj(a0) + proj(a1) = W(a0 + a1) + n_nbr·bias ; then gate ; then mva_proj + residual
                ops.gemm(ws.ao[rows], cv.proj.weight, cv.proj.bias, ws.xm[rows], geo=geo, bias_scale=n_nbr,
                         gate=(tab(mod_mva), 2))
                ops.gemm(ws.xm[rows], self.mva_proj.weight, self.mva_proj.bias, x[rows], geo=geo, res=x[rows])

            ctx.spatial_attention(ops, ws.qkv, ws.ao, d, ctx.kv_map, scale, cv_qkv_rows, cv_proj_rows)

        # (g) MLP
        ops.ln_modulate(x, ws.xm, mod, 3, 4, geo)
        ops.gemm(ws.xm, W(self.mlp.fc1), self.mlp.fc1.bias, ws.hid, geo=geo, act="gelu")
        ops.gemm(ws.hid, W(self.mlp.fc2), self.mlp.fc2.bias, x, geo=geo, gate=(mod, 5), res=x)

        # (h) control skip: x_base += after_proj(x_control)
        if self.is_control_block and skip_into is not None:
            ops.gemm(x, self.after_proj.weight, self.after_proj.bias, skip_into, geo=geo, res=skip_into)
