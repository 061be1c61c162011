"""Condition encoders of the STDiT3 drop-in.

These produce the *timestep-independent* operands of the hot path — the condition sequence
``y[β,T,L,C]`` and the BEV-map tokens — and therefore run ONCE per sample, outside the 30-step
denoising loop (the drop-in caches their outputs, see ``stdit3.MagicDriveSTDiT3``).  They account for
<1 % of a step's FLOPs and are evaluated with PyTorch library ops (cuBLAS/cuDNN); the per-step work
(blocks, patch-embed, final layer) goes through the sm_100a kernels in ``csrc/``.

Parameter names/shapes follow the reference so its checkpoints load with ``strict=True``:
``magicdrivedit/models/magicdrive/embedder.py`` (bbox :354-484, cam :535-675, map :678-833) and
``magicdrivedit/models/layers/blocks.py`` (timestep :913-952, size :985-1016, caption :1019-1064).
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn
import torch.nn.functional as F


# ----------------------------------------------------------------------------------------------
# small building blocks
# ----------------------------------------------------------------------------------------------
class RMSNormWeight(nn.Module):
    """Holder of the per-head RMSNorm gain (``LlamaRMSNorm``, blocks.py:54-68); eps 1e-6."""

    def __init__(self, dim: int, eps: float = 1e-6):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))
        self.eps = eps

    def forward(self, x):
        xf = x.float()
        xf = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + self.eps)
        return self.weight * xf.to(x.dtype)


class RotaryFreqs(nn.Module):
    """Holds ``freqs`` like rotary-embedding-torch 0.5.3 (key ``….rope.freqs``); the parameter
    follows the model dtype, so angles are bf16(pos)·bf16(freq) in a bf16 model."""

    def __init__(self, dim: int, theta: float = 10000.0):
        super().__init__()
        self.freqs = nn.Parameter(
            1.0 / (theta ** (torch.arange(0, dim, 2)[: dim // 2].float() / dim)), requires_grad=False)

    def angles(self, T: int, dtype, device):
        """[T, dim] pair-wise duplicated angles, computed in ``dtype`` like the reference."""
        pos = torch.arange(T, device=device, dtype=dtype).to(self.freqs.dtype)
        ang = pos[:, None] * self.freqs[None, :]
        return ang.repeat_interleave(2, dim=-1)

    def rotate(self, t):
        """t: [..., T, D]; interleaved-pair rotation."""
        ang = self.angles(t.shape[-2], t.dtype, t.device)
        x1, x2 = t[..., 0::2], t[..., 1::2]
        rot = torch.stack((-x2, x1), dim=-1).flatten(-2)
        return t * ang.cos() + rot * ang.sin()


class MlpParams(nn.Module):
    """fc1 -> GELU(tanh) -> fc2 (timm 0.9.16 ``Mlp`` key layout)."""

    def __init__(self, in_features, hidden_features, out_features=None):
        super().__init__()
        self.fc1 = nn.Linear(in_features, hidden_features)
        self.fc2 = nn.Linear(hidden_features, out_features or in_features)

    def forward(self, x):
        return self.fc2(F.gelu(self.fc1(x), approximate="tanh"))


class SelfAttnParams(nn.Module):
    """qkv / q_norm / k_norm / proj holder (``Attention`` and ``MultiHeadAttention`` key layout)."""

    def __init__(self, dim, num_heads, qkv_bias=True, qk_norm=True):
        super().__init__()
        self.dim, self.num_heads, self.head_dim = dim, num_heads, dim // num_heads
        self.qkv = nn.Linear(dim, 3 * dim, bias=qkv_bias)
        self.q_norm = RMSNormWeight(self.head_dim) if qk_norm else nn.Identity()
        self.k_norm = RMSNormWeight(self.head_dim) if qk_norm else nn.Identity()
        self.proj = nn.Linear(dim, dim)


class CrossAttnParams(nn.Module):
    """q_linear / kv_linear / proj holder (``MultiHeadCrossAttention`` key layout)."""

    def __init__(self, dim, num_heads):
        super().__init__()
        self.dim, self.num_heads, self.head_dim = dim, num_heads, dim // num_heads
        self.q_linear = nn.Linear(dim, dim)
        self.kv_linear = nn.Linear(dim, 2 * dim)
        self.proj = nn.Linear(dim, dim)


def _softmax_attention(q, k, v, scale):
    """Eager attention with the reference's rounding points (blocks.py:257-267)."""
    attn = ((q * scale) @ k.transpose(-2, -1)).float().softmax(dim=-1).to(q.dtype)
    return attn @ v


def cog_temp_down(x):
    """[B,T,N,D] -> keep frame 0, average pairs (1,2),(3,4),… (embedder.py:117-141); even T: plain
    pair-average."""
    B, T, N, D = x.shape
    if T % 2 == 1:
        if T == 1:
            return x
        rest = x[:, 1:].reshape(B, (T - 1) // 2, 2, N, D).mean(dim=2)
        return torch.cat([x[:, :1], rest], dim=1)
    return x.reshape(B, T // 2, 2, N, D).mean(dim=2)


def _make_downsampler(factor):
    if factor == -1:
        return lambda x: x.mean(dim=1, keepdim=True)
    if factor == 4.5:
        return lambda x: cog_temp_down(cog_temp_down(x))
    if factor == 0:
        return lambda x: x
    raise NotImplementedError(f"time_downsample_factor={factor}")


class _TemporalLayer(nn.Module):
    """One pre-LN transformer layer over the frame axis with RoPE, shared by the box and the
    ego-motion ("frame") embedders (embedder.py:397-483, :605-675)."""

    def _init_temporal(self, hidden, num_heads, mlp_ratio, qk_norm, use_table, after_proj, tdf):
        self.hidden_size = hidden
        self.rope = RotaryFreqs(hidden // num_heads)
        self.attn = SelfAttnParams(hidden, num_heads, qkv_bias=True, qk_norm=qk_norm)
        self.scale_shift_table = nn.Parameter(torch.randn(6, hidden) / hidden ** 0.5) if use_table else None
        self.mlp = MlpParams(hidden, int(hidden * mlp_ratio))
        self.final_proj = nn.Linear(hidden, hidden) if after_proj else None
        if self.final_proj is not None:
            nn.init.zeros_(self.final_proj.weight)
            nn.init.zeros_(self.final_proj.bias)
        self.downsampler = _make_downsampler(tdf)

    def _temporal(self, x):
        """x: [(b n), T, d]."""
        if self.scale_shift_table is not None:
            sh1, sc1, g1, sh2, sc2, g2 = self.scale_shift_table[None].chunk(6, dim=1)
        else:
            sh1 = sh2 = sc1 = sc2 = 0.0
            g1 = g2 = 1.0
        a = self.attn
        Bn, T, d = x.shape
        h = F.layer_norm(x, (d,), eps=1e-6) * (1 + sc1) + sh1
        qkv = a.qkv(h).view(Bn, T, 3, a.num_heads, a.head_dim).permute(2, 0, 3, 1, 4)
        q, k, v = qkv.unbind(0)
        q, k = a.q_norm(q), a.k_norm(k)
        q, k = self.rope.rotate(q), self.rope.rotate(k)
        o = _softmax_attention(q, k, v, a.head_dim ** -0.5).transpose(1, 2).reshape(Bn, T, d)
        x = x + g1 * a.proj(o)
        h = F.layer_norm(x, (d,), eps=1e-6) * (1 + sc2) + sh2
        return x + g2 * self.mlp(h)


def _fourier(x, num_freqs):
    """[x, sin(x·2^0), cos(x·2^0), …, sin(x·2^(F-1)), cos(x·2^(F-1))] on the last dim
    (embedder.py:487-532 with include_input, log_sampling)."""
    outs = [x]
    for i in range(num_freqs):
        f = 2.0 ** i
        outs += [torch.sin(x * f), torch.cos(x * f)]
    return torch.cat(outs, dim=-1)


# ----------------------------------------------------------------------------------------------
# timestep / fps / caption
# ----------------------------------------------------------------------------------------------
def sinusoid(t, dim=256, max_period=10000):
    """[cos | sin] embedding in fp32 (blocks.py:926-946)."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(half, dtype=torch.float32, device=t.device) / half)
    args = t[:, None].float() * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


class TimestepEmbedder(nn.Module):
    def __init__(self, hidden_size, frequency_embedding_size=256):
        super().__init__()
        self.mlp = nn.Sequential(nn.Linear(frequency_embedding_size, hidden_size), nn.SiLU(),
                                 nn.Linear(hidden_size, hidden_size))
        self.frequency_embedding_size = frequency_embedding_size

    def forward(self, t, dtype):
        return self.mlp(sinusoid(t, self.frequency_embedding_size).to(dtype))


class SizeEmbedder(TimestepEmbedder):
    def __init__(self, hidden_size, frequency_embedding_size=256):
        super().__init__(hidden_size, frequency_embedding_size)
        self.outdim = hidden_size

    def forward(self, s, bs):
        if s.ndim == 1:
            s = s[:, None]
        if s.shape[0] != bs:
            s = s.repeat(bs // s.shape[0], 1)
        b, dims = s.shape
        dtype = self.mlp[0].weight.dtype
        emb = self.mlp(sinusoid(s.reshape(-1), self.frequency_embedding_size).to(dtype))
        return emb.reshape(b, dims * self.outdim)


class CaptionEmbedder(nn.Module):
    def __init__(self, in_channels, hidden_size, uncond_prob, token_num=120):
        super().__init__()
        self.y_proj = MlpParams(in_channels, hidden_size, hidden_size)
        self.register_buffer("y_embedding", torch.randn(token_num, in_channels) / in_channels ** 0.5)
        self.uncond_prob = uncond_prob

    def forward(self, caption, train=False, force_drop_ids=None):
        if force_drop_ids is not None:
            drop = force_drop_ids == 1
            caption = torch.where(drop[:, None, None, None], self.y_embedding, caption)
        return self.y_proj(caption)


# ----------------------------------------------------------------------------------------------
# camera / ego-motion
# ----------------------------------------------------------------------------------------------
class CamEmbedder(nn.Module):
    def __init__(self, input_dim, out_dim, num=7, num_freqs=4, include_input=True, log_sampling=True,
                 after_proj=False, **_):
        super().__init__()
        assert include_input and log_sampling
        self.num_freqs = num_freqs
        self.emb2token = nn.Linear(input_dim * (1 + 2 * num_freqs) * num, out_dim)
        self.uncond_cam = nn.Parameter(torch.randn(input_dim, num))
        self.after_proj = nn.Linear(out_dim, out_dim) if after_proj else None
        if self.after_proj is not None:
            nn.init.zeros_(self.after_proj.weight)
            nn.init.zeros_(self.after_proj.bias)

    def _tokens(self, param, mask):
        if param.shape[1] == 4:
            param = param[:, :-1]
        bs = param.shape[0]
        if mask is not None:
            param = torch.where((mask > 0)[:, None, None], param, self.uncond_cam[None])
        emb = _fourier(param.transpose(1, 2).reshape(-1, param.shape[1]), self.num_freqs)
        return self.emb2token(emb.reshape(bs, -1))

    def embed_cam(self, param, mask=None, **kwargs):
        token = self._tokens(param, mask)
        if self.after_proj is not None:
            token = self.after_proj(token)
        return token, None


class CamEmbedderTemp(CamEmbedder, _TemporalLayer):
    def __init__(self, input_dim, out_dim, num=7, num_freqs=4, include_input=True, log_sampling=True,
                 after_proj=False, num_heads=8, mlp_ratio=4.0, qk_norm=False,
                 use_scale_shift_table=False, time_downsample_factor=-1, **_):
        CamEmbedder.__init__(self, input_dim, out_dim, num, num_freqs, include_input, log_sampling,
                             after_proj=False)
        self._init_temporal(out_dim, num_heads, mlp_ratio, qk_norm, use_scale_shift_table,
                            after_proj, time_downsample_factor)

    def embed_cam(self, param, mask=None, T=None, S=None):
        token = self._tokens(param, mask)                      # [(b T S), d]
        d = token.shape[-1]
        x = token.view(-1, T, S, d).transpose(1, 2).reshape(-1, T, d)
        x = self._temporal(x)
        x = x.view(-1, S, T, d).transpose(1, 2)                # [b, T, S, d]
        if self.final_proj is not None:
            x = self.final_proj(x)
        return self.downsampler(x), None


# ----------------------------------------------------------------------------------------------
# 3-D boxes
# ----------------------------------------------------------------------------------------------
_XYZ_MIN = (-200.0, -300.0, -20.0)
_XYZ_RANGE = (350.0, 650.0, 80.0)


class ContinuousBBoxWithTextTempEmbedding(nn.Module, ):
    NuScenes_bbox_classes = ["car", "truck", "construction_vehicle", "bus", "trailer", "barrier", "motorcycle",
                             "bicycle", "pedestrian", "traffic_cone"]

    def __init__(self, n_classes, class_token_dim=768, trainable_class_token=False, embedder_num_freq=4,
                 proj_dims=(768, 512, 512, 768), mode="cxyz", minmax_normalize=True,
                 use_text_encoder_init=True, after_proj=False, sample_id=False, num_heads=8,
                 mlp_ratio=4.0, qk_norm=False, use_scale_shift_table=False,
                 time_downsample_factor=-1, **_):
        super().__init__()
        n_pts = {"cxyz": 4, "all-xyz": 8}[mode]
        self.mode, self.minmax_normalize, self.num_freqs = mode, minmax_normalize, embedder_num_freq
        self.use_text_encoder_init = use_text_encoder_init
        pos_dim = 3 * (1 + 2 * embedder_num_freq) * n_pts
        self.bbox_proj = nn.Linear(pos_dim, proj_dims[0])
        self.second_linear = nn.Sequential(
            nn.Linear(proj_dims[0] + class_token_dim, proj_dims[1]), nn.SiLU(),
            nn.Linear(proj_dims[1], proj_dims[2]), nn.SiLU(),
            nn.Linear(proj_dims[2], proj_dims[3]))
        tokens = torch.randn(n_classes, class_token_dim)
        if trainable_class_token:
            self._class_tokens = nn.Parameter(tokens)
        else:
            self.register_buffer("_class_tokens", tokens)
        self.mean_var = nn.Parameter(torch.randn(n_classes, 2)) if sample_id else None
        self.null_class_feature = nn.Parameter(torch.zeros(class_token_dim))
        self.null_pos_feature = nn.Parameter(torch.zeros(pos_dim))
        self.mask_class_feature = nn.Parameter(torch.zeros(class_token_dim))
        self.mask_pos_feature = nn.Parameter(torch.zeros(pos_dim))
        _TemporalLayer._init_temporal(self, proj_dims[-1], num_heads, mlp_ratio, qk_norm,
                                      use_scale_shift_table, after_proj, time_downsample_factor)

    _temporal = _TemporalLayer._temporal

    @property
    def class_tokens(self):
        return self._class_tokens

    @property
    def box_latent_shape(self):
        return self._class_tokens[0:1].shape

    def prepare(self, text_encoder, classes=None):
        """embedder.py:253-261: class tokens from the text encoder (mean over its tokens) when so configured."""
        if self.use_text_encoder_init:
            self.set_category_token(text_encoder, self.NuScenes_bbox_classes if classes is None else classes)

    @torch.no_grad()
    def set_category_token(self, text_encoder, class_names):
        for idx, name in enumerate(class_names):                 # embedder.py:263-271
            self._class_tokens[idx].copy_(text_encoder([name]).mean(dim=1)[0])

    def forward_feature(self, pos_emb, cls_emb):
        emb = F.silu(self.bbox_proj(pos_emb))
        return self.second_linear(torch.cat([emb, cls_emb], dim=-1))

    def forward(self, bboxes, classes, null_mask=None, mask=None, box_latent=None, **_):
        """bboxes [B,T,N,P,3]; classes/null_mask/mask [B,T,N]; -> [B, T', N, d]."""
        B, T, N = classes.shape
        ref = self.null_pos_feature
        pts = bboxes.reshape(B * T * N, *bboxes.shape[3:])

        def flat(m):
            m = torch.ones(B * T * N, device=pts.device) if m is None else m.reshape(-1)
            return m[:, None].to(ref.dtype)

        mask, null_mask = flat(mask), flat(null_mask)
        if self.minmax_normalize:
            lo = torch.as_tensor(_XYZ_MIN, dtype=pts.dtype, device=pts.device)
            rg = torch.as_tensor(_XYZ_RANGE, dtype=pts.dtype, device=pts.device)
            pts = (pts - lo) / rg
        pos = _fourier(pts, self.num_freqs).reshape(pts.shape[0], -1).to(ref.dtype)
        pos = pos * null_mask + self.null_pos_feature[None] * (1 - null_mask)
        pos = pos * mask + self.mask_pos_feature[None] * (1 - mask)

        idx = classes.reshape(-1).long()
        cls = self._class_tokens[idx]
        if self.mean_var is not None:
            mu, logvar = self.mean_var[idx].split(1, dim=1)
            if box_latent is None:
                box_latent = torch.randn_like(cls)
            else:
                box_latent = box_latent.reshape(cls.shape)
            cls = cls + (box_latent * torch.exp(0.5 * logvar) + mu)
        cls = cls * null_mask + self.null_class_feature[None] * (1 - null_mask)
        cls = cls * mask + self.mask_class_feature[None] * (1 - mask)

        emb = self.forward_feature(pos, cls)                    # [(B T N), d]
        d = emb.shape[-1]
        x = emb.view(B, T, N, d).permute(0, 2, 1, 3).reshape(B * N, T, d)
        x = self._temporal(x)
        x = x.view(B, N, T, d).transpose(1, 2)                  # [B, T, N, d]
        if self.final_proj is not None:
            x = self.final_proj(x)
        return self.downsampler(x)


# ----------------------------------------------------------------------------------------------
# BEV map
# ----------------------------------------------------------------------------------------------
class MapControlEmbedding(nn.Module):
    def __init__(self, conditioning_embedding_channels=320, conditioning_size=(25, 200, 200),
                 block_out_channels=(32, 64, 128, 256), use_uncond_map=None, drop_cond_ratio=0.0):
        super().__init__()
        self.uncond_map = None
        ch = list(block_out_channels)
        self.conv_in = nn.Conv2d(conditioning_size[0], ch[0], kernel_size=3, padding=1)
        blocks = []
        for i in range(len(ch) - 2):
            blocks.append(nn.Conv2d(ch[i], ch[i], kernel_size=3, padding=1))
            blocks.append(nn.Conv2d(ch[i], ch[i + 1], kernel_size=3, padding=(2, 1), stride=2))
        blocks.append(nn.Conv2d(ch[-2], ch[-2], kernel_size=3, padding=(2, 1)))
        blocks.append(nn.Conv2d(ch[-2], ch[-1], kernel_size=3, padding=(2, 1), stride=(2, 1)))
        self.blocks = nn.ModuleList(blocks)
        self.conv_out = nn.Conv2d(ch[-1], conditioning_embedding_channels, kernel_size=3, padding=1)

    def forward(self, cond):
        h = F.silu(self.conv_in(cond))
        for blk in self.blocks:
            h = F.silu(blk(h))
        return self.conv_out(h)


class _CogDown3D(nn.Module):
    """``CogVideoXDownsample3D`` with compress_time (embedder.py:41-95); key ``conv.*``."""

    def __init__(self, cin, cout):
        super().__init__()
        self.conv = nn.Conv2d(cin, cout, kernel_size=3, stride=1, padding=0)

    def forward(self, x):                                       # [b, C, T, H, W]
        b, C, T, H, W = x.shape
        x = cog_temp_down(x.permute(0, 2, 3, 4, 1).reshape(b, T, H * W, C))
        T2 = x.shape[1]
        x = x.reshape(b * T2, H, W, C).permute(0, 3, 1, 2)
        x = self.conv(F.pad(x, (0, 1, 0, 1)))
        return x.reshape(b, T2, *x.shape[1:]).permute(0, 2, 1, 3, 4)


class MapControlTempEmbedding(nn.Module):
    def __init__(self, hidden_size, time_downsample_factor):
        super().__init__()
        if time_downsample_factor != 4.5:
            raise NotImplementedError("only the CogVideoX-style (4.5) temporal down-sampler is built")
        # indices 0/2 are parameter-free ZeroPad2d in the reference -> keys conv_blocks.{1,3}.conv.*
        self.conv_blocks = nn.Sequential(
            nn.ZeroPad2d((1, 0, 1, 0)), _CogDown3D(hidden_size // 2, hidden_size // 2),
            nn.ZeroPad2d((1, 0, 1, 0)), _CogDown3D(hidden_size // 2, hidden_size))

    def forward(self, x):
        return self.conv_blocks(x)


EMBEDDER_CLASSES = {
    "CamEmbedder": CamEmbedder,
    "CamEmbedderTemp": CamEmbedderTemp,
    "ContinuousBBoxWithTextTempEmbedding": ContinuousBBoxWithTextTempEmbedding,
    "MapControlEmbedding": MapControlEmbedding,
}


def resolve_embedder(dotted: str):
    """The reference names embedder classes by dotted string (``load_module``, utils.py:22-27);
    the last component selects the native class."""
    name = dotted.rsplit(".", 1)[-1]
    if name not in EMBEDDER_CLASSES:
        raise KeyError(f"unsupported embedder class {dotted}")
    return EMBEDDER_CLASSES[name]
