"""``MagicDriveSTDiT3`` — B200-native drop-in for the registry-built denoiser of the reference
(``magicdrivedit/models/magicdrive/magicdrive_stdit3.py:475-1244``).

Same constructor kwargs, same ``forward(x, timestep, y, maps, bbox, cams, rel_pos, fps, height,
width, …)`` signature, same state-dict key layout, fp32 ``[b, C·NC, T, H, W]`` output.  The per-step
work (patch-embed, the 82 blocks, final layer, unpatchify) is executed by the hand-written sm_100a
library behind ``ops`` on one canonical activation layout ``[β, T, S, C]``; there is no eager
PyTorch path for it and the module refuses to run when the CUDA library is missing.
"""
from __future__ import annotations

import contextlib
import logging
import math
import os

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
from transformers import PreTrainedModel

from . import indexmaps
from .blocks import Dims, STDiT3Block, Workspace, W
from .config import MagicDriveSTDiT3Config
from .embedders import (CaptionEmbedder, MapControlTempEmbedding, RotaryFreqs, SizeEmbedder,
                        TimestepEmbedder, resolve_embedder)
from .registry import MODELS

logger = logging.getLogger(__name__)


class PatchEmbedParams(nn.Module):
    """Holder of a ``PatchEmbed3D`` conv kernel (blocks.py:101-155): key ``proj.{weight,bias}`` with
    weight ``[C, Cin, 1, 2, 2]``; applied as a per-patch linear over ``Cin·4`` gathered inputs."""

    def __init__(self, patch_size, in_chans, embed_dim):
        super().__init__()
        self.patch_size = tuple(patch_size)
        self.in_chans, self.embed_dim = in_chans, embed_dim
        self.proj = nn.Conv3d(in_chans, embed_dim, kernel_size=self.patch_size, stride=self.patch_size)

    def weight2d(self):
        return self.proj.weight.view(self.embed_dim, -1)


class PositionEmbedding2D(nn.Module):
    """2-D sin/cos position table (blocks.py:1067-1116): [sin(w f), cos(w f), sin(h f), cos(h f)]."""

    def __init__(self, dim: int):
        super().__init__()
        assert dim % 4 == 0, "dim must be divisible by 4"
        self.dim = dim
        half = dim // 2
        self.register_buffer("inv_freq", 1.0 / (10000 ** (torch.arange(0, half, 2).float() / half)),
                             persistent=False)
        self._cache = {}

    def forward(self, device, dtype, h, w, scale=1.0, base_size=None):
        key = (str(device), dtype, h, w, scale, base_size)
        if key not in self._cache:
            gh = torch.arange(h, device=device) / scale
            gw = torch.arange(w, device=device) / scale
            if base_size is not None:
                gh = gh * (base_size / h)
                gw = gw * (base_size / w)
            # token n = h'·W' + w'
            tok_h = gh[:, None].expand(h, w).reshape(-1)
            tok_w = gw[None, :].expand(h, w).reshape(-1)
            inv = self.inv_freq.to(device=device, dtype=torch.float32)

            def sc(t):
                o = t[:, None] * inv[None, :]
                return torch.cat((torch.sin(o), torch.cos(o)), dim=-1)
            # the reference names the w-grid "grid_h" (meshgrid called with w first): w comes first
            self._cache[key] = torch.cat([sc(tok_w), sc(tok_h)], dim=-1).to(dtype).contiguous()
        return self._cache[key]                                  # [S, C]


class FinalLayerParams(nn.Module):
    """``T2IFinalLayer`` parameters (blocks.py:870-905)."""

    def __init__(self, hidden_size, num_patch, out_channels):
        super().__init__()
        self.linear = nn.Linear(hidden_size, num_patch * out_channels)
        self.scale_shift_table = nn.Parameter(torch.randn(2, hidden_size) / hidden_size ** 0.5)
        self.out_channels = out_channels


class _Ctx:
    """Forward-wide constants handed to every block."""
    pass


class MagicDriveSTDiT3(PreTrainedModel):
    """See module docstring.  Attribute surface used by callers is preserved: ``y_embedder``
    (+ ``y_embedding``), ``camera_embedder.uncond_cam``, ``frame_embedder.uncond_cam``,
    ``sample_box_latent``, ``device``, ``in_channels``, ``config``."""

    config_class = MagicDriveSTDiT3Config
    _ops = None           # ops backend (the sm_100a library binding); tests may inject a spec backend
    _BLOCK_LISTS = ("base_blocks_s", "base_blocks_t", "control_blocks_s", "control_blocks_t",
                    "brushnet_blocks_s", "brushnet_blocks_t")     # (the last two: BrushNet variant, brushnet.py)

    def __init__(self, config: MagicDriveSTDiT3Config):
        super().__init__(config)
        c = config
        assert not c.use_st_cross_attn, "STDiT3 have temporal downsample, this means nothing."
        assert not c.sequence_parallelism_temporal, "not support!"
        self.pred_sigma = c.pred_sigma
        self.in_channels = c.in_channels
        self.out_channels = c.in_channels * 2 if c.pred_sigma else c.in_channels
        self.depth, self.control_depth = c.depth, c.control_depth
        self.mlp_ratio, self.hidden_size, self.num_heads = c.mlp_ratio, c.hidden_size, c.num_heads
        self.enable_sequence_parallelism = c.enable_sequence_parallelism
        self.patch_size = tuple(c.patch_size)
        assert self.patch_size == (1, 2, 2), "the native patch kernels are built for patch (1,2,2)"
        self.input_sq_size = c.input_sq_size
        self.force_pad_h_for_sp_size = c.force_pad_h_for_sp_size
        self.micro_frame_size = c.micro_frame_size
        C, NH = self.hidden_size, self.num_heads

        self.pos_embed = PositionEmbedding2D(C)
        self.rope = RotaryFreqs(C // NH)
        self.x_embedder = PatchEmbedParams(self.patch_size, self.in_channels, C)
        self.t_embedder = TimestepEmbedder(C)
        self.t_block = nn.Sequential(nn.SiLU(), nn.Linear(C, 6 * C))
        self.y_embedder = CaptionEmbedder(c.caption_channels, C, c.class_dropout_prob, c.model_max_length)
        self.fps_embedder = SizeEmbedder(C)
        self.x_control_embedder = (PatchEmbedParams(self.patch_size, self.in_channels, C)
                                   if c.use_x_control_embedder else None)
        self.register_buffer("base_token", torch.randn(C))
        self.camera_embedder = resolve_embedder(c.cam_encoder_cls)(out_dim=C, **c.cam_encoder_param)
        self.frame_embedder = resolve_embedder(c.frame_emb_cls)(out_dim=C, **c.frame_emb_param)
        self.bbox_embedder = resolve_embedder(c.bbox_embedder_cls)(**c.bbox_embedder_param)
        self.controlnet_cond_embedder = resolve_embedder(c.map_embedder_cls)(
            conditioning_embedding_channels=C // 2, **c.map_embedder_param)
        self.controlnet_cond_embedder_temp = MapControlTempEmbedding(C, c.map_embedder_downsample_rate)
        self.controlnet_cond_patchifier = PatchEmbedParams(self.patch_size, C, C)

        def blocks(n, **kw):
            return nn.ModuleList([STDiT3Block(C, NH, self.mlp_ratio, qk_norm=c.qk_norm, **kw)
                                  for _ in range(n)])

        self.base_blocks_s = blocks(self.depth)
        self.base_blocks_t = blocks(self.depth, temporal=True) if c.with_temp_block else None
        self.before_proj = nn.Linear(C, C)
        self.control_blocks_s = blocks(self.control_depth, is_control_block=True,
                                       skip_cross_view=c.control_skip_cross_view)
        self.control_blocks_t = (None if c.control_skip_temporal else
                                 blocks(self.control_depth, temporal=True, is_control_block=True))
        self.final_layer = FinalLayerParams(C, int(np.prod(self.patch_size)), self.out_channels)
        self.initialize_weights()
        self._cond_cache = None
        self._ws = None
        self.cache_conditions = True
        self.sp_group = None          # set by sp.set_sequence_parallel_group or explicitly

    # ------------------------------------------------------------------------------------------
    def initialize_weights(self):
        """Init recipe of magicdrive_stdit3.py:778-836 (xavier linears, zero-init new branches)."""
        def _basic(m):
            if isinstance(m, nn.Linear):
                nn.init.xavier_uniform_(m.weight)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
        self.apply(_basic)

        def _zero(m):
            if isinstance(m, nn.Linear):
                nn.init.zeros_(m.weight)
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
        for blk in self.base_blocks_s:
            _zero(blk.mva_proj)
        if self.base_blocks_t is not None:
            for blk in self.base_blocks_t:
                _zero(blk.attn.proj)
                _zero(blk.cross_attn.proj)      # (mlp.fc2 is NOT zeroed upstream: :807 is a no-op)
        for blk in self.control_blocks_s:
            _zero(blk.mva_proj)
            _zero(blk.after_proj)
        if self.control_blocks_t is not None:
            for blk in self.control_blocks_t:
                _zero(blk.after_proj)
        _zero(self.before_proj)
        _zero(self.bbox_embedder.final_proj)
        _zero(self.camera_embedder.after_proj)
        _zero(self.frame_embedder.final_proj)
        w = self.controlnet_cond_patchifier.proj.weight.data
        nn.init.xavier_uniform_(w.view(w.shape[0], -1))
        for m in (self.bbox_embedder.mlp, self.frame_embedder.mlp):
            nn.init.normal_(m.fc1.weight, std=0.02)
            nn.init.normal_(m.fc2.weight, std=0.02)
        nn.init.normal_(self.camera_embedder.emb2token.weight, std=0.02)

    def _init_weights(self, module):      # transformers hook; real init is initialize_weights()
        return

    def get_dynamic_size(self, x):
        _, _, T, H, W = x.size()
        return indexmaps.dynamic_size(T, H, W, self.patch_size)

    def sample_box_latent(self, n_boxes, generator=None):
        if self.bbox_embedder.mean_var is None:
            return None
        return torch.randn((n_boxes, self.bbox_embedder.box_latent_shape[1]), generator=generator)

    # ------------------------------------------------------------------------------------------
    @property
    def ops(self):
        if MagicDriveSTDiT3._ops is None:
            from . import ops as _native          # raises if libmdv2.so is missing / no CUDA
            MagicDriveSTDiT3._ops = _native
        return MagicDriveSTDiT3._ops

    def clear_condition_cache(self):
        self._cond_cache = None

    def _apply(self, fn, *a, **k):
        self._cond_cache = None
        self._ws = None
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, *a, **k):
        self._cond_cache = None
        return super().load_state_dict(*a, **k)

    # ------------------------------------------------------------------------------------------
    # condition encoding (timestep-independent; SURVEY §8 a6/a7, cached across denoising steps)
    # ------------------------------------------------------------------------------------------
    def encode_text(self, y, mask=None, drop_cond_mask=None):
        """magicdrive_stdit3.py:860-880: caption MLP, clipped to max(mask.sum)+1 tokens."""
        force = None if drop_cond_mask is None else 1 - drop_cond_mask
        y = self.y_embedder(y, False, force_drop_ids=force)
        if mask is not None:
            if mask.shape[0] != y.shape[0]:
                mask = mask.repeat(y.shape[0] // mask.shape[0], 1)
            y_lens = [i + 1 for i in mask.sum(dim=1).tolist()]
            max_len = int(min(max(y_lens), y.shape[2]))
            if drop_cond_mask is not None and not drop_cond_mask.all():        # on any drop this must be the maximum
                assert max_len == y.shape[2]
            y = y.squeeze(1)[:, :max_len]
        else:
            y = y.squeeze(1)
        return y

    def prepare_text_embedding(self, text_encoder):
        """magicdrive_stdit3.py:1009-1019: class tokens of the box embedder and ``base_token`` from the text encoder
        (``text -> caption MLP -> the prompt's tokens``); invalidates the condition cache."""
        @torch.no_grad()
        def text_to_embedding(text):
            ret = text_encoder.encode(text)
            hidden_state = self.encode_text(ret["y"], mask=None)
            return hidden_state[:, :int(ret["mask"].sum(dim=1))]
        self.bbox_embedder.prepare(text_to_embedding)
        with torch.no_grad():
            self.base_token[:] = text_to_embedding("").squeeze()
        self.clear_condition_cache()

    def encode_cond_sequence(self, bbox, cams, rel_pos, y, mask, drop_cond_mask, drop_frame_mask):
        """-> [β, T', L, C] = frame(1) ‖ cam(1) ‖ text ‖ boxes (magicdrive_stdit3.py:918-961)."""
        b = len(y)
        NC = cams.shape[0] // b
        y = self.encode_text(y, mask, drop_cond_mask).repeat_interleave(NC, dim=0)
        parts = []
        if bbox is not None:
            keep = torch.logical_and(drop_cond_mask[:, None], drop_frame_mask).repeat_interleave(NC, dim=0)
            m = bbox["masks"]
            keep = keep[:, :, None].expand_as(m)
            null_mask = torch.ones_like(m)
            null_mask[m == 0] = 0
            vis = torch.ones_like(m)
            vis[m == -1] = 0
            vis[torch.logical_and(m == 1, keep == 0)] = 0
            emb = self.bbox_embedder(bboxes=bbox["bboxes"], classes=bbox["classes"].to(torch.int32),
                                     null_mask=null_mask, mask=vis, box_latent=bbox.get("box_latent"))
            parts.append(self.base_token[None, None, None] + emb)

        def cam_tokens(cam, embedder, drop):                     # cam [β, T, S, 3|4, n]
            B_, T_, S_ = cam.shape[:3]
            mk = drop.repeat_interleave(NC, dim=0)[:, :, None].expand(B_, T_, S_).reshape(-1)
            tok, _ = embedder.embed_cam(cam.reshape(B_ * T_ * S_, *cam.shape[3:]), mk, T=T_, S=S_)
            return tok
        cam_emb = cam_tokens(cams[:, 0:1], self.camera_embedder, drop_cond_mask[:, None])
        cam_emb = cam_emb.view(cams.shape[0], 1, cams.shape[2], -1)
        frame_emb = self.base_token[None, None, None] + cam_tokens(rel_pos, self.frame_embedder, drop_frame_mask)
        cam_emb = self.base_token[None, None, None] + cam_emb
        Tc = frame_emb.shape[1]
        cam_emb = cam_emb.expand(-1, Tc, -1, -1)
        y = y[:, None].expand(-1, Tc, -1, -1)
        return torch.cat([frame_emb, cam_emb, y] + parts, dim=2)

    def encode_map_features(self, maps, x_shape_thw, h_pad):
        """BEV-map CNN + temporal down-sampler + nearest resize to the latent grid + H-pad
        (magicdrive_stdit3.py:963-1004) -> [b, C, T, Hx(+pad), Wx]; patchify/GEMMs follow natively."""
        b, Tp = maps.shape[:2]
        f = self.controlnet_cond_embedder(maps.flatten(0, 1))
        f = f.view(b, Tp, *f.shape[1:]).permute(0, 2, 1, 3, 4)
        if self.micro_frame_size is None:
            f = self.controlnet_cond_embedder_temp(f)
        else:
            f = torch.cat([self.controlnet_cond_embedder_temp(f[:, :, i:i + self.micro_frame_size])
                           for i in range(0, f.shape[2], self.micro_frame_size)], dim=2)
        if tuple(f.shape[-3:]) != tuple(x_shape_thw):
            f = F.interpolate(f, tuple(x_shape_thw))
        if h_pad > 0:
            f = F.pad(f, (0, 0, 0, h_pad * self.patch_size[1]))
        return f.contiguous()

    def _all_blocks(self):
        """(name, block) for every block that owns a cross_attn, in a fixed order."""
        out = []
        for name in self._BLOCK_LISTS:
            lst = getattr(self, name, None)
            if lst is not None:
                out += [(f"{name}.{i}", blk) for i, blk in enumerate(lst)]
        return out

    def _condition_key(self, tensors):
        key = []
        for t in tensors:
            if isinstance(t, torch.Tensor):
                key.append((id(t), t._version, tuple(t.shape), t.data_ptr()))
            else:
                key.append(t)
        return tuple(key)

    @torch.no_grad()
    def _conditions(self, ops, dtype, y, mask, bbox, cams, rel_pos, maps, drop_cond_mask,
                    drop_frame_mask, geom, drop_key=None):
        """Returns the cached timestep-independent operands:
        cond [β,Tc,L,C], per-block cross-attn K/V, map tokens ``m`` [b·T·S, C] (before_proj applied)."""
        (b, NC, T, Hp, Wp, h_pad, x_thw) = geom
        # (drop masks the caller left at None are materialised anew by every forward: key them as None, not by identity)
        dk = drop_key if drop_key is not None else (drop_cond_mask, drop_frame_mask)
        flat = [y, mask, cams, rel_pos, maps, dk[0], dk[1], geom, dtype]
        if bbox is not None:
            flat += [bbox[k] for k in sorted(bbox)]
        key = self._condition_key(flat)
        # two entries: the "slice" samplers alternate between the conditional and the null condition set
        for entry in (self._cond_cache or ()) if self.cache_conditions else ():
            if entry[0] == key:
                return entry[1]
        C = self.hidden_size

        def cast(t):     # callers hand conditions in the model dtype (inference_magicdrive.py:345-358); be lenient
            return t.to(dtype) if isinstance(t, torch.Tensor) and t.is_floating_point() else t
        y, cams, rel_pos, maps = cast(y), cast(cams), cast(rel_pos), cast(maps)
        drop_cond_mask, drop_frame_mask = cast(drop_cond_mask), cast(drop_frame_mask)
        if bbox is not None:
            bbox = {k_: cast(v_) for k_, v_ in bbox.items()}
        tf32 = contextlib.nullcontext()
        if dtype == torch.float32 and y.is_cuda:
            tf32 = torch.backends.cudnn.flags(enabled=True, allow_tf32=False)
        with tf32:
            cond = self.encode_cond_sequence(bbox, cams, rel_pos, y, mask, drop_cond_mask, drop_frame_mask)
            if cond.shape[1] != T and cond.shape[1] > 1:
                L_ = cond.shape[2]
                c2 = cond.permute(0, 2, 3, 1).reshape(cond.shape[0], L_ * C, cond.shape[1])
                cond = F.interpolate(c2, T).reshape(cond.shape[0], L_, C, T).permute(0, 3, 1, 2)
            cond = cond.to(dtype).contiguous()
            feat = self.encode_map_features(maps, x_thw, h_pad).to(dtype)
        Bc, Tc, L, _ = cond.shape
        cond2d = cond.view(Bc * Tc * L, C)
        y_kv = {}
        for name, blk in self._all_blocks():
            if not blk.skip_cross_attn:
                kv = torch.empty(cond2d.shape[0], 2 * C, dtype=dtype, device=cond.device)
                ops.gemm(cond2d, W(blk.cross_attn.kv_linear), blk.cross_attn.kv_linear.bias, kv)
                y_kv[name] = kv
        # map tokens: patchify -> patchifier GEMM -> before_proj GEMM   (rows: b, T, S)
        S = Hp * Wp
        n_rows = b * T * S
        patches = torch.empty(n_rows, 4 * C, dtype=dtype, device=cond.device)
        ops.patchify(feat, patches, 1, Hp, Wp, 0, S)
        m0 = torch.empty(n_rows, C, dtype=dtype, device=cond.device)
        pf = self.controlnet_cond_patchifier
        ops.gemm(patches, pf.weight2d(), pf.proj.bias, m0)
        m = torch.empty(n_rows, C, dtype=dtype, device=cond.device)
        ops.gemm(m0, self.before_proj.weight, self.before_proj.bias, m)
        out = dict(cond=cond, y_kv=y_kv, m=m.view(b, T, S, C), L=L, y_frames=Tc)
        if self.cache_conditions:
            # the key holds ids / data pointers: keep the keyed tensors alive so neither can be recycled
            self._cond_cache = [(key, out, flat)] + list(self._cond_cache or ())[:1]
        return out

    # ------------------------------------------------------------------------------------------
    # the denoising step
    # ------------------------------------------------------------------------------------------
    def _stacked_tables(self, names, attr, dtype):
        """The ``attr`` parameters of every block of the lists ``names``, stacked ``[n_total, K, C]`` in the model dtype
        — built once and kept until a parameter changes — plus the slice of every list."""
        lists = [(n, getattr(self, n)) for n in names if getattr(self, n, None) is not None]
        params = [getattr(blk, attr) for _, blocks in lists for blk in blocks]
        key = (dtype, tuple((p_.data_ptr(), p_._version) for p_ in params))
        cache = self.__dict__.setdefault("_tab_cache", {})
        hit = cache.get((names, attr))
        if hit is None or hit[0] != key:
            stacked = torch.stack([p_.detach() for p_ in params]).to(dtype).contiguous()
            slices, o = {}, 0
            for n, blocks in lists:
                slices[n] = slice(o, o + len(blocks))
                o += len(blocks)
            hit = cache[(names, attr)] = (key, stacked, slices)
        return hit[1], hit[2]

    def _mod_tables(self, names, attr, tsel, dtype):
        """{list name: [n_blk, b·T, K, C] fp32} = table + (x_mask-selected) timestep vector, rounded like the
        reference's bf16 add (magicdrive_stdit3.py:213-221) — ONE libmdv2 launch for all the lists ``names``.
        tsel: [b, T, K, C] (or [b, T, K, 1])."""
        stacked, slices = self._stacked_tables(names, attr, dtype)
        b, T, K = tsel.shape[:3]
        mod = self.ops.adaln_tables(stacked, tsel.to(dtype).reshape(b * T, K, tsel.shape[-1]).contiguous())
        return {n: mod[sl] for n, sl in slices.items()}

    @torch.no_grad()
    def forward(self, x, timestep, y, maps, bbox, cams, rel_pos, fps, height, width,
                drop_cond_mask=None, drop_frame_mask=None, mv_order_map=None, t_order_map=None,
                mask=None, x_mask=None, **kwargs):
        ops = self.ops
        inpaint = kwargs.pop("_inpaint", None)             # (x_inpaint, mask_inpaint): BrushNet variant only
        sde = kwargs.pop("_sde", None)                     # (timestep_inpaint, num_timesteps, noise): SDE variant only
        dtype = self.x_embedder.proj.weight.dtype
        dev = x.device
        b, real_T = x.size(0), rel_pos.size(1)
        drop_key = (drop_cond_mask, drop_frame_mask)
        if drop_cond_mask is None:
            drop_cond_mask = torch.ones((b,), device=dev, dtype=x.dtype)
        if drop_frame_mask is None:
            drop_frame_mask = torch.ones((b, real_T), device=dev, dtype=x.dtype)
        assert mv_order_map is not None
        NC = len(mv_order_map)
        C, NH = self.hidden_size, self.num_heads
        x = x.to(dtype).contiguous()
        timestep = timestep.to(dtype)
        # (y keeps the caller's tensor: the condition cache is keyed on its identity — a cast here would make a new
        # tensor, and a cache miss, every step whenever the text encoder's dtype differs from the model's; the cast
        # happens inside _conditions, next to cams / rel_pos / maps)
        assert x.shape[1] == self.in_channels * NC

        # --- geometry (P2/P3) ---
        _, _, Tx, Hx, Wx = x.shape
        T, Hp, Wp = indexmaps.dynamic_size(Tx, Hx, Wx, self.patch_size)
        sp_group, p, rank = self._sp()
        q = self.force_pad_h_for_sp_size if self.force_pad_h_for_sp_size is not None else (p if p > 1 else None)
        h_pad = indexmaps.h_pad_for(Hp, Wp, q)
        Hp += h_pad
        S = Hp * Wp
        if p > 1:
            assert S % p == 0, f"S={S} should be divisible by {p}!"
            assert NH % p == 0, f"num_heads={NH} should be divisible by {p}!"
        s0, s1 = indexmaps.sp_slice(S, rank, p)
        Sl = s1 - s0

        base_size = round(S ** 0.5)
        # (the reference reads height[0] / width[0] on the host every forward, magicdrive_stdit3.py:1113; the device ->
        # host read is done once per pair of caller tensors, so a sampling loop — and a captured CUDA graph — has none)
        hw_key = (id(height), height._version, id(width), width._version)
        if getattr(self, "_hw_scale", (None,))[0] != hw_key:
            self._hw_scale = (hw_key, (float(height[0]) * float(width[0])) ** 0.5 / self.input_sq_size, height, width)
        scale = self._hw_scale[1]
        pos = self.pos_embed(dev, dtype, Hp, Wp, scale=scale, base_size=base_size)       # [S, C]

        # --- timestep embeddings (b rows; library GEMVs) ---
        fps_e = self.fps_embedder(fps.unsqueeze(1), b)
        t = self.t_embedder(timestep, dtype=dtype) + fps_e
        t_mlp = self.t_block(t)
        if x_mask is not None:
            t0 = self.t_embedder(torch.zeros_like(timestep), dtype=dtype) + fps_e
            t0_mlp = self.t_block(t0)
            sel = x_mask.to(dev).bool()[:, :, None]                                    # [b, T, 1]
            tm_sel = torch.where(sel, t_mlp[:, None], t0_mlp[:, None])                 # [b, T, 6C]
        else:
            t0 = t0_mlp = None
            tm_sel = t_mlp[:, None].expand(b, T, 6 * C)
        tm_sel_brush = tm_sel
        if sde is not None:                                # the third stream is modulated by its own time vector
            tc, tc0 = self._sde_time_vectors(sde[0], dtype, fps_e, t_mlp, t0, t0_mlp)
            tm_sel_brush = (torch.where(sel, tc[:, None], tc0[:, None]) if x_mask is not None
                            else tc[:, None].expand(b, T, 6 * C))

        # --- conditions (cached) ---
        geom = (b, NC, T, Hp, Wp, h_pad, (Tx, Hx, Wx))
        cnd = self._conditions(ops, dtype, y, mask, bbox, cams, rel_pos, maps, drop_cond_mask,
                               drop_frame_mask, geom, drop_key=drop_key)

        # --- embed x: two streams share one patch gather ---
        d = Dims(b=b, NC=NC, T=T, S=Sl, C=C, H=NH, D=C // NH)
        if self._ws is None or self._ws.key != (d.N, C, dtype, dev, self.mlp_ratio):
            self._ws = Workspace(d, dtype, dev, self.mlp_ratio)
            self._ws.ao_sym = None
            from . import sp
            if p > 1 and sp.peer_exchange_available(x, sp_group):
                # the attention output buffer is written by the peers (NVLink scatter): symmetric allocation
                self._ws.ao_sym = sp.symm_buffer("ao", (d.N, C), self._ws.ao, sp_group)
                self._ws.ao = self._ws.ao_sym[0]
        ws = self._ws
        N = d.N
        patches = torch.empty(N, self.in_channels * 4, dtype=dtype, device=dev)
        ops.patchify(x, patches, NC, Hp, Wp, s0, Sl)
        pos_l = pos[s0:s1].contiguous()
        xb = torch.empty(N, C, dtype=dtype, device=dev)
        ops.gemm(patches, self.x_embedder.weight2d(), self.x_embedder.proj.bias, xb,
                 res=pos_l, res_map=(Sl, 0))
        xe = self.x_control_embedder if self.x_control_embedder is not None else self.x_embedder
        pm = (pos_l[None, None] + cnd["m"][:, :, s0:s1]).reshape(b * T * Sl, C)          # [b·T·Sl, C]
        xc = torch.empty(N, C, dtype=dtype, device=dev)
        ops.gemm(patches, xe.weight2d(), xe.proj.bias, xc, res=pm, res_map=(T * Sl, NC))

        # third stream (BrushNet): tokens of [x | encoded masked video | mask], one skip per block into xb
        xi = None
        if inpaint is not None:
            xi = self._embed_inpaint(ops, dtype, x, inpaint, (b, NC, T, Tx, Hx, Wx, Hp, Wp, s0, Sl), pos_l, sde)

        # --- modulation tables for all blocks at once ---
        tm6 = tm_sel.reshape(b, T, 6, C)
        tm3 = tm_sel[:, :, :3].reshape(b, T, 3, 1)
        tm6b = tm_sel_brush.reshape(b, T, 6, C)
        main_lists = tuple(n for n in self._BLOCK_LISTS if not n.startswith("brushnet"))
        mods = self._mod_tables(main_lists, "scale_shift_table", tm6, dtype)
        mva_lists = ("base_blocks_s",) + (("control_blocks_s",) if self.control_depth > 0 and
                                          not self.control_blocks_s[0].skip_cross_view else ())
        mva = self._mod_tables(mva_lists, "scale_shift_table_mva", tm3, dtype)
        mod_mva, mva_ctrl, mva_brush = mva["base_blocks_s"], mva.get("control_blocks_s"), None
        if xi is not None:                                 # the third stream has its own time vector (SDE variant)
            mods.update(self._mod_tables(tuple(n for n in self._BLOCK_LISTS if n.startswith("brushnet")),
                                         "scale_shift_table", tm6b, dtype))
            if not self.brushnet_blocks_s[0].skip_cross_view:
                mva_brush = self._mod_tables(("brushnet_blocks_s",), "scale_shift_table_mva",
                                             tm_sel_brush[:, :, :3].reshape(b, T, 3, 1), dtype)["brushnet_blocks_s"]

        ctx = _Ctx()
        ctx.L, ctx.y_frames = cnd["L"], cnd["y_frames"]
        # index map and RoPE tables: small, shape-keyed, kept on the device (no host -> device copy per step)
        step_key = (tuple((k_, tuple(int(n_) for n_ in v_)) for k_, v_ in sorted(mv_order_map.items())), b, T, NC,
                    dtype, str(dev), self.rope.freqs.data_ptr(), self.rope.freqs._version)
        if getattr(self, "_step_tabs", (None,))[0] != step_key:
            ang = self.rope.angles(T, dtype, dev)                                      # [T, D]
            self._step_tabs = (step_key, indexmaps.cross_view_kv_map(mv_order_map, b, T, NC).to(dev),
                               ang.cos().float().contiguous(), ang.sin().float().contiguous())
        _, ctx.kv_map, ctx.rope_cos, ctx.rope_sin = self._step_tabs
        ctx.spatial_attention = self._spatial_attention
        ykv = cnd["y_kv"]

        # --- blocks (magicdrive_stdit3.py:1173-1198) ---
        for i in range(self.depth):
            self.base_blocks_s[i].run(ops, xb, ws, d, mods["base_blocks_s"][i], mod_mva[i],
                                      ykv.get(f"base_blocks_s.{i}"), ctx)
            if i < self.control_depth:
                self.control_blocks_s[i].run(ops, xc, ws, d, mods["control_blocks_s"][i],
                                             None if mva_ctrl is None else mva_ctrl[i],
                                             ykv.get(f"control_blocks_s.{i}"), ctx, skip_into=xb)
            if xi is not None:                             # magicdrive_stdit3.py:1714-1728 / :1790-1804
                self.brushnet_blocks_s[i].run(ops, xi, ws, d, mods["brushnet_blocks_s"][i],
                                              None if mva_brush is None else mva_brush[i],
                                              ykv.get(f"brushnet_blocks_s.{i}"), ctx, skip_into=xb)
            if self.base_blocks_t is not None:
                self.base_blocks_t[i].run(ops, xb, ws, d, mods["base_blocks_t"][i], None,
                                          ykv.get(f"base_blocks_t.{i}"), ctx)
            if i < self.control_depth and self.control_blocks_t is not None:
                self.control_blocks_t[i].run(ops, xc, ws, d, mods["control_blocks_t"][i], None,
                                             ykv.get(f"control_blocks_t.{i}"), ctx, skip_into=xb)
            # (reference: below control_depth the brushnet temporal block runs whenever it exists,
            #  magicdrive_stdit3.py:1766-1781; above it only together with the base temporal blocks, :1812-1842)
            if xi is not None and self.brushnet_blocks_t is not None and (
                    i < self.control_depth or self.base_blocks_t is not None):
                self.brushnet_blocks_t[i].run(ops, xi, ws, d, mods["brushnet_blocks_t"][i], None,
                                              ykv.get(f"brushnet_blocks_t.{i}"), ctx, skip_into=xb)

        # --- final layer (blocks.py:893-905) + unpatchify (P7) ---
        fl = self.final_layer
        geo = (T, Sl, NC)

        def final_mod(tv):                                                             # tv [b, C]
            m_ = (fl.scale_shift_table.to(dtype)[None] + tv[:, None].to(dtype)).float()  # [b, 2, C]
            return m_[:, None].expand(b, T, 2, C).reshape(b * T, 2, C).contiguous()
        ops.ln_modulate(xb, ws.xm, final_mod(t), 0, 1, geo)
        if x_mask is not None:
            # reference quirk (blocks.py:899-903): the t0 branch normalises the ALREADY modulated
            # tensor, i.e. masked frames get mod_t0(LN(mod_t(LN(x)))); replicated, not "fixed".
            ops.ln_modulate(ws.xm, ws.ao, final_mod(t0), 0, 1, geo)
            keep = x_mask.to(dev).bool().repeat_interleave(NC, dim=0)[:, :, None, None]   # [β,T,1,1]
            ws.xm.view(d.B, T, Sl, C).copy_(torch.where(keep, ws.xm.view(d.B, T, Sl, C),
                                                        ws.ao.view(d.B, T, Sl, C)))
        n_out = fl.linear.out_features
        yo = torch.empty(N, n_out, dtype=dtype, device=dev)
        ops.gemm(ws.xm, fl.linear.weight, fl.linear.bias, yo)
        if p > 1:
            yo = self._sp_gather_tokens(yo, d, p, sp_group)                              # [β·T·S, n_out]
        out = torch.empty(b, self.out_channels * NC, Tx, Hx, Wx, dtype=torch.float32, device=dev)
        ops.unpatchify(yo, out, NC, T, Hp, Wp)
        return out

    # ------------------------------------------------------------------------------------------
    # sequence parallelism (Ulysses head-scatter over the spatial tokens; blocks.py:546-645)
    # ------------------------------------------------------------------------------------------
    def capture_step(self, warmup=2, **inputs):
        """-> ``GraphedStep``: this forward, with these (fixed) conditions, as a replayable CUDA graph."""
        return GraphedStep(self, inputs, warmup=warmup)

    def _sp(self):
        from . import sp
        if not self.enable_sequence_parallelism:
            return None, 1, 0
        group = self.sp_group if self.sp_group is not None else sp.get_sequence_parallel_group()
        return group, torch.distributed.get_world_size(group), torch.distributed.get_rank(group)

    def _spatial_attention(self, ops, qkv, out, d: Dims, kv_map, scale, produce, consume):
        """Attention over the S tokens of every (β, t) frame (qkv: [N, 3C], columns q|k|v); with kv_map
        the K/V come from the neighbour views and the per-neighbour results are summed (cross-view, P5).

        ``produce(rows, tab)`` fills ``qkv[rows]`` (LN-modulate + qkv projection) and ``consume(rows, tab)``
        folds ``out[rows]`` back into the residual stream (output projection); ``tab(mod)`` narrows a
        modulation table to the scheduler-batch entries of those rows.  On one GPU both run once over all
        rows; under sequence parallelism they run per chunk so the all-to-alls hide behind them."""
        group, p, _ = self._sp()
        if p == 1:
            C = d.C
            rows = slice(0, d.N)
            produce(rows, lambda m: m)
            ops.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], out, d.B * d.T, d.S, d.S, d.H, d.D, scale,
                          kv_map=kv_map)
            consume(rows, lambda m: m)
            return
        from . import sp
        if self._ws is not None and getattr(self._ws, "ao_sym", None) is not None:
            _, _, rank = self._sp()
            sp.ulysses_attention_peer(ops, qkv, self._ws.ao_sym, d, kv_map, scale, group, p, rank, produce, consume)
            return
        sp.ulysses_attention(ops, qkv, out, d, kv_map, scale, group, p, produce, consume)

    def _sp_gather_tokens(self, yo, d: Dims, p, group):
        from . import sp
        return sp.gather_tokens(yo, d, p, group)


class GraphedStep:
    """One denoising step of a sample as a CUDA graph (the reference launches ~4000 eager kernels per
    step, schedulers/rf/__init__.py:82-113).  The conditions are fixed for the sample (encoded once, before the capture,
    through the condition cache); every call refreshes the latent and the timestep in place and replays the ~1100
    libmdv2 launches of the step without touching Python or the launch path.

        step = model.capture_step(**inputs)          # inputs: the forward's keyword arguments, device tensors
        v = step(x, timestep)                        # -> the same fp32 velocity tensor object every call
    """

    def __init__(self, model, inputs, warmup=2):
        assert model.cache_conditions, "a captured step needs the condition cache (conditions are encoded once)"
        self.model = model
        self.static = dict(inputs)
        self.static["x"] = inputs["x"].clone()
        self.static["timestep"] = inputs["timestep"].clone()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                      # warm-up off the capture: workspaces, kernel attributes,
            for _ in range(warmup):                        # condition cache, symmetric buffers
                model(**self.static)
        torch.cuda.current_stream().wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = model(**self.static)

    def __call__(self, x, timestep):
        self.static["x"].copy_(x)
        self.static["timestep"].copy_(timestep)
        self.graph.replay()
        return self.out


def _build(cls, from_pretrained=None, force_huggingface=False, strict=True, **kwargs):
    """Body shared by the three registry factories (magicdrive_stdit3.py:2554-2620)."""
    if from_pretrained is not None and not os.path.exists(from_pretrained):
        return cls.from_pretrained(from_pretrained, **kwargs)
    from_pretrained_pixart = kwargs.pop("from_pretrained_pixart", None)
    config = MagicDriveSTDiT3Config(depth=28, hidden_size=1152, patch_size=(1, 2, 2), num_heads=16, **kwargs)
    model = cls(config)
    from . import checkpoint
    if from_pretrained is not None and force_huggingface:       # initialise from an STDiT3 base checkpoint
        checkpoint.load_from_stdit3_pretrained(model, from_pretrained)
    elif from_pretrained is not None:
        checkpoint.load_checkpoint(model, from_pretrained, strict=strict)
    elif from_pretrained_pixart is not None:
        checkpoint.load_from_pixart_pretrained(model, from_pretrained_pixart)
    else:
        logger.info("Your model does not use any pre-trained model.")
    return model


@MODELS.register_module("MagicDriveSTDiT3-XL/2")
def MagicDriveSTDiT3_XL_2(from_pretrained=None, force_huggingface=False, **kwargs):
    """Registry factory with the reference's name and signature (magicdrive_stdit3.py:2554-2570)."""
    return _build(MagicDriveSTDiT3, from_pretrained, force_huggingface, **kwargs)
