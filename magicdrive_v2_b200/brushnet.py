"""BrushNet variant of the denoiser: a third stream of ``depth`` control-style blocks fed with
``[x | shallow-encoded masked video | mask]`` whose per-block ``after_proj`` skips are added to the base stream.

Mirror of ``MagicDriveSTDiT3BrushNet`` and ``ShallowEncoder`` (``magicdrive_stdit3.py:1247-1867``): same constructor
(config keys ``brushnet_skip_cross_attn``, ``control_skip_cross_view``), same
``forward(x, x_inpaint, mask_inpaint, timestep, y, maps, bbox, cams, rel_pos, fps, height, width, ...)`` signature,
same state-dict keys (``shallow_encoder.*``, ``brushnet_blocks_s/t.*``, ``x_brushnet_embedder.*``).  The blocks run on
the same libmdv2 kernels as the base and control streams; the pixel-space ``ShallowEncoder`` convolutions and the
trilinear mask resize do not depend on the timestep — they run once per sample on stock torch ops and are cached
across the sampling steps like the other condition encoders.  LoRA adapters on the base blocks
(``use_lora_base_blocks``, ``:1422-1464``) keep their ``lora_A`` / ``lora_B`` state-dict entries and are folded into
the projection weights once per load (``W + α/r · B·A``), so the hot path runs the same GEMMs.

``MagicDriveSTDiT3SDEBrushNet`` (``magicdrive_stdit3.py:1870-2470``) gives the third stream its own noise level: the
shallow-encoded video is mixed with a noise latent at ``timestep_inpaint`` (``add_noise``, ``:1976-1994``) and the
brushnet blocks are modulated by ``t_combine_block(cat[t_block(t), t_inpaint_block(t_inpaint)])`` (``:2215-2232``).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from .blocks import STDiT3Block
from .config import MagicDriveSTDiT3Config
from .registry import MODELS
from .stdit3 import MagicDriveSTDiT3, PatchEmbedParams, _build


_LORA_TARGETS = (("attn", ("qkv", "proj")), ("cross_attn", ("q_linear", "kv_linear", "proj")), ("mlp", ("fc1", "fc2")))


def _add_lora(lin: nn.Linear, rank: int, alpha: float) -> nn.Linear:
    """``_add_lora_to_linear`` (magicdrive_stdit3.py:1422-1464): ``lora_A [r, in]`` (kaiming), ``lora_B [out, r]``
    (zeros) registered on the Linear itself, so the keys are ``….qkv.lora_A`` / ``….qkv.lora_B``."""
    lin.lora_A = nn.Parameter(torch.zeros(rank, lin.in_features))
    lin.lora_B = nn.Parameter(torch.zeros(lin.out_features, rank))
    nn.init.kaiming_uniform_(lin.lora_A, a=math.sqrt(5))
    lin.lora_scaling = alpha / rank
    lin._w_eff = None
    lin._w_key = None
    return lin


class ShallowEncoder(nn.Module):
    """(β, 3, 4t+1, 8h, 8w) -> (β, out_channels, t, h, w): three stride-2 convs + one conv per frame, then a
    (5,1,1)/stride-4 temporal conv (magicdrive_stdit3.py:1247-1311)."""

    def __init__(self, in_channels=3, out_channels=4, temporal_downsample=4):
        super().__init__()
        self.temporal_downsample = temporal_downsample
        self.conv1 = nn.Conv2d(in_channels, 64, kernel_size=3, stride=2, padding=1)
        self.conv2 = nn.Conv2d(64, 128, kernel_size=3, stride=2, padding=1)
        self.conv3 = nn.Conv2d(128, 256, kernel_size=3, stride=2, padding=1)
        self.conv4 = nn.Conv2d(256, out_channels, kernel_size=3, stride=1, padding=1)
        td = temporal_downsample
        self.temporal_conv = nn.Conv3d(out_channels, out_channels, kernel_size=(td + 1, 1, 1), stride=(td, 1, 1),
                                       padding=(td // 2, 0, 0))

    def forward(self, x, chunk_size=64):
        B, _, T, _, _ = x.shape
        x = x.permute(0, 2, 1, 3, 4).flatten(0, 1)                                   # (β t) c h w
        chunks = []
        for i in range(0, x.shape[0], chunk_size):
            ch = F.relu(self.conv1(x[i:i + chunk_size]))
            ch = F.relu(self.conv2(ch))
            ch = F.relu(self.conv3(ch))
            chunks.append(self.conv4(ch))
        x = torch.cat(chunks, dim=0)
        x = x.view(B, T, *x.shape[1:]).permute(0, 2, 1, 3, 4)
        return self.temporal_conv(x)


class MagicDriveSTDiT3BrushNet(MagicDriveSTDiT3):
    def __init__(self, config: MagicDriveSTDiT3Config):
        super().__init__(config)
        c = config
        if getattr(c, "lora_dropout", 0.0):
            raise NotImplementedError("LoRA dropout is a training-time feature; the inference drop-in folds the adapters")
        C, NH = self.hidden_size, self.num_heads
        self.shallow_encoder = ShallowEncoder(3, self.in_channels, temporal_downsample=4)
        skip_ca = getattr(c, "brushnet_skip_cross_attn", True)
        self.brushnet_blocks_s = nn.ModuleList([
            STDiT3Block(C, NH, self.mlp_ratio, qk_norm=c.qk_norm, is_control_block=True, skip_cross_attn=skip_ca,
                        skip_cross_view=c.control_skip_cross_view) for _ in range(self.depth)])
        self.brushnet_blocks_t = nn.ModuleList([
            STDiT3Block(C, NH, self.mlp_ratio, qk_norm=c.qk_norm, temporal=True, is_control_block=True,
                        skip_cross_attn=skip_ca) for _ in range(self.depth)])
        self.x_brushnet_embedder = PatchEmbedParams(self.patch_size, self.in_channels * 2 + 1, C)
        self._inpaint_cache = None
        self._lora = []
        if getattr(c, "use_lora_base_blocks", False):       # magicdrive_stdit3.py:1374-1378, :1398-1420
            for lst in (self.base_blocks_s, self.base_blocks_t):
                for blk in (lst if lst is not None else []):
                    for sub, names in _LORA_TARGETS:
                        for n in names:
                            lin = getattr(getattr(blk, sub, None), n, None)
                            if isinstance(lin, nn.Linear):
                                _add_lora(lin, c.lora_rank, c.lora_alpha)
                                # reference quirk, replicated: the spatial MultiHeadAttention applies its qkv with
                                # F.linear on weight slices (blocks.py:438-445), which by-passes the patched forward —
                                # those adapters exist in the state dict but never act
                                if not (sub == "attn" and n == "qkv" and not blk.temporal):
                                    self._lora.append(lin)

    def _drop_folded(self):
        for lin in self._lora:
            lin._w_eff = None
            lin._w_key = None

    def _apply(self, fn, *a, **k):
        self._inpaint_cache = None
        self._drop_folded()
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, *a, **k):
        self._drop_folded()
        return super().load_state_dict(*a, **k)

    @torch.no_grad()
    def _fold_lora(self):
        """W_eff = W + (alpha / rank) · B·A, in fp32, rounded once to the model dtype; refreshed when a parameter
        was written since the last fold."""
        for lin in self._lora:
            key = tuple((t.data_ptr(), t._version) for t in (lin.weight, lin.lora_A, lin.lora_B))
            if lin._w_eff is None or lin._w_key != key:
                w = lin.weight.float() + lin.lora_scaling * (lin.lora_B.float() @ lin.lora_A.float())
                lin._w_eff, lin._w_key = w.to(lin.weight.dtype).contiguous(), key
                self._cond_cache = None                     # cached cross-attention K/V used the old kv_linear

    @torch.no_grad()
    def _encode_inpaint(self, dtype, x_inpaint, mask_inpaint, NC, thw):
        """Timestep-independent half of the third stream: -> (encoded video [β, Cin, T, Hx, Wx], mask [β, 1, T, Hx, Wx]).
        Cached on the identity of the two inputs (the sampler hands the same tensors every step)."""
        key = self._condition_key([x_inpaint, mask_inpaint, NC, tuple(thw), dtype])
        if self.cache_conditions and self._inpaint_cache is not None and self._inpaint_cache[0] == key:
            return self._inpaint_cache[1]

        def views(t):                                       # "B (C NC) T ... -> (B NC) C T ..."
            B_, CN = t.shape[:2]
            return t.to(dtype).view(B_, CN // NC, NC, *t.shape[2:]).transpose(1, 2).flatten(0, 1)
        enc = self.shallow_encoder(views(x_inpaint))
        mk = F.interpolate(views(mask_inpaint), size=tuple(thw), mode="trilinear", align_corners=False)
        assert tuple(enc.shape[-3:]) == tuple(thw) == tuple(mk.shape[-3:]), \
            f"After encoding - x:{tuple(thw)}, x_inpaint_encoded:{tuple(enc.shape)}, mask_inpaint:{tuple(mk.shape)}"
        out = (enc.contiguous(), mk.contiguous())
        if self.cache_conditions:
            self._inpaint_cache = (key, out, (x_inpaint, mask_inpaint))
        return out

    def _embed_inpaint(self, ops, dtype, x, inpaint, geometry, pos_l, sde=None):
        """x_brushnet_embedder(cat[x, enc, mask]) + pos  ->  [N, C] tokens of the third stream
        (magicdrive_stdit3.py:1648-1654); rows in the canonical (β, t, s) order, local S-slice under SP."""
        b, NC, T, Tx, Hx, Wx, Hp, Wp, s0, Sl = geometry
        x_inpaint, mask_inpaint = inpaint
        enc, mk = self._encode_inpaint(dtype, x_inpaint, mask_inpaint, NC, (Tx, Hx, Wx))
        if sde is not None:
            enc = self._noise_inpaint(dtype, enc, NC, *sde)
        xv = x.view(b, self.in_channels, NC, Tx, Hx, Wx).transpose(1, 2).flatten(0, 1)   # (b NC) C T H W
        # 2·Cin + 1 = 33 channels -> 132 gathered inputs per patch; zero channels pad that to a multiple of 64 so the
        # embedding runs on the tcgen05 GEMM like the other two streams
        cin = 2 * self.in_channels + 1
        cpad = -(-cin * 4 // 64) * 16
        zeros = torch.zeros(xv.shape[0], cpad - cin, *xv.shape[2:], dtype=dtype, device=x.device)
        xcat = torch.cat([xv, enc, mk, zeros], dim=1).contiguous()
        n_rows = b * NC * T * Sl
        patches = torch.empty(n_rows, cpad * 4, dtype=dtype, device=x.device)
        ops.patchify(xcat, patches, 1, Hp, Wp, s0, Sl)
        w = self.x_brushnet_embedder.proj.weight                                      # [C, cin, 1, 2, 2]
        w2d = F.pad(w.view(w.shape[0], cin, 4), (0, 0, 0, cpad - cin)).reshape(w.shape[0], cpad * 4).contiguous()
        xi = torch.empty(n_rows, self.hidden_size, dtype=dtype, device=x.device)
        ops.gemm(patches, w2d, self.x_brushnet_embedder.proj.bias, xi, res=pos_l, res_map=(Sl, 0))
        return xi

    @torch.no_grad()
    def forward(self, x, x_inpaint, mask_inpaint, timestep, y, maps, bbox, cams, rel_pos, fps, height, width,
                drop_cond_mask=None, drop_frame_mask=None, mv_order_map=None, t_order_map=None, mask=None,
                x_mask=None, **kwargs):
        self._fold_lora()
        return super().forward(x, timestep, y, maps, bbox, cams, rel_pos, fps, height, width,
                               drop_cond_mask=drop_cond_mask, drop_frame_mask=drop_frame_mask,
                               mv_order_map=mv_order_map, t_order_map=t_order_map, mask=mask, x_mask=x_mask,
                               _inpaint=(x_inpaint, mask_inpaint), **kwargs)


class MagicDriveSTDiT3SDEBrushNet(MagicDriveSTDiT3BrushNet):
    """BrushNet whose inpaint branch runs at an independent timestep (magicdrive_stdit3.py:1870-2470).  Same modules as
    the BrushNet variant plus ``t_inpaint_block`` and ``t_combine_block``; forward takes ``timestep_inpaint``,
    ``num_timesteps`` and ``noise_inpaint_encoded`` (``:1996-2020``)."""

    def __init__(self, config: MagicDriveSTDiT3Config):
        super().__init__(config)
        C = self.hidden_size
        self.t_inpaint_block = nn.Sequential(nn.SiLU(), nn.Linear(C, 6 * C, bias=True))
        self.t_combine_block = nn.Sequential(nn.SiLU(), nn.Linear(12 * C, 6 * C, bias=True))

    @staticmethod
    def add_noise(original_samples, noise, timesteps, num_timesteps):
        """``timepoint·x + (1 - timepoint)·noise`` with ``timepoint = 1 - t / num_timesteps`` per sample, in fp32
        (magicdrive_stdit3.py:1976-1994)."""
        tp = 1 - timesteps.float() / num_timesteps
        tp = tp.view(-1, 1, 1, 1, 1)
        return tp * original_samples + (1 - tp) * noise

    def _noise_inpaint(self, dtype, enc, NC, timestep_inpaint, num_timesteps, noise):
        """Noised encoding of the masked video (magicdrive_stdit3.py:2099-2111); the clean encoding stays cached, the
        mix depends on ``timestep_inpaint`` and is redone every step."""
        if noise is None:
            # the reference draws phase-preserving "structured" noise on the CPU with the un-vendored
            # ``structured_noise`` package (:2057-2096); its arithmetic is not on this box, so the caller supplies it
            raise NotImplementedError("noise_inpaint_encoded=None needs the third-party structured_noise package; "
                                      "pass noise_inpaint_encoded [b, C·NC, T, h, w]")
        B_, CN = noise.shape[:2]
        nz = noise.to(dtype).view(B_, CN // NC, NC, *noise.shape[2:]).transpose(1, 2).flatten(0, 1)
        assert nz.shape == enc.shape, \
            f"noise_inpaint_encoded shape {tuple(nz.shape)} does not match x_inpaint_encoded shape {tuple(enc.shape)}!"
        ts = timestep_inpaint.to(enc.device).repeat_interleave(NC, dim=0)
        return self.add_noise(enc, nz, ts, num_timesteps).to(dtype).contiguous()

    def _sde_time_vectors(self, timestep_inpaint, dtype, fps_e, t_mlp, t0, t0_mlp):
        """-> (t_combined_mlp, t0_combined_mlp | None), each [b, 6C] (magicdrive_stdit3.py:2215-2232).  Unlike
        ``timestep``, ``timestep_inpaint`` is NOT rounded to the model dtype before the sinusoid (reference quirk)."""
        t_in = self.t_embedder(timestep_inpaint.to(fps_e.device), dtype=dtype) + fps_e
        tc = self.t_combine_block(torch.cat([t_mlp, self.t_inpaint_block(t_in)], dim=-1))
        tc0 = None
        if t0 is not None:
            tc0 = self.t_combine_block(torch.cat([t0_mlp, self.t_inpaint_block(t0)], dim=-1))
        return tc, tc0

    @torch.no_grad()
    def forward(self, x, x_inpaint, mask_inpaint, timestep, timestep_inpaint, num_timesteps, y, maps, bbox, cams,
                rel_pos, fps, height, width, noise_inpaint_encoded=None, drop_cond_mask=None, drop_frame_mask=None,
                mv_order_map=None, t_order_map=None, mask=None, x_mask=None, **kwargs):
        self._fold_lora()
        return MagicDriveSTDiT3.forward(
            self, x, timestep, y, maps, bbox, cams, rel_pos, fps, height, width, drop_cond_mask=drop_cond_mask,
            drop_frame_mask=drop_frame_mask, mv_order_map=mv_order_map, t_order_map=t_order_map, mask=mask,
            x_mask=x_mask, _inpaint=(x_inpaint, mask_inpaint),
            _sde=(timestep_inpaint, num_timesteps, noise_inpaint_encoded), **kwargs)


@MODELS.register_module("MagicDriveSTDiT3-XL/2-BrushNet")
def MagicDriveSTDiT3_XL_2_BrushNet(from_pretrained=None, force_huggingface=False, **kwargs):
    """Registry factory with the reference's name and signature (magicdrive_stdit3.py:2573-2595); like the reference
    it loads non-strictly, so a base-model checkpoint initialises the shared streams."""
    return _build(MagicDriveSTDiT3BrushNet, from_pretrained, force_huggingface, strict=False, **kwargs)


@MODELS.register_module("MagicDriveSTDiT3-XL/2-SDEBrushNet")
def MagicDriveSTDiT3_XL_2_SDEBrushNet(from_pretrained=None, force_huggingface=False, **kwargs):
    """Registry factory with the reference's name and signature (magicdrive_stdit3.py:2598-2620)."""
    return _build(MagicDriveSTDiT3SDEBrushNet, from_pretrained, force_huggingface, strict=False, **kwargs)
