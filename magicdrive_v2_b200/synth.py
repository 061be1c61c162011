"""Synthetic workload definitions for the STDiT3 denoise step.

Constructor kwargs mirror ``configs/magicdrive/test/17-16x848x1600_…fsp8_cfg2.0.py:76-150`` of the
reference; inputs mirror the tensors ``scripts/inference_magicdrive.py:345-463`` hands to
``model.forward``.  Everything is seeded with ``torch.Generator`` on the CPU so that the oracle,
the golden fixtures and the CUDA path see identical bytes.
"""
from __future__ import annotations

import math

import torch

_EMB = "magicdrivedit.models.magicdrive.embedder."

MV_ORDER_MAP = {0: [5, 1], 1: [0, 2], 2: [1, 3], 3: [2, 4], 4: [3, 5], 5: [4, 0]}


def model_kwargs(hidden_size=1152, depth=28, num_heads=16, control_depth=13, *,
                 video=True, force_pad_h_for_sp_size=8, caption_channels=4096,
                 model_max_length=300, in_channels=16, map_size=400, embed_heads=8,
                 enable_sequence_parallelism=False):
    """Reference constructor kwargs.  ``video=False`` gives the stage-1 image config
    (``inference/1x224x400_…noTemp…py``): no temporal blocks, pooled conditions."""
    tdf = 4.5 if video else -1
    temp_kw = dict(num_heads=embed_heads, mlp_ratio=4.0, qk_norm=True, enable_flash_attn=False,
                   enable_xformers=False, enable_layernorm_kernel=False,
                   use_scale_shift_table=True, time_downsample_factor=tdf)
    return dict(
        depth=depth, hidden_size=hidden_size, patch_size=(1, 2, 2), num_heads=num_heads,
        in_channels=in_channels, caption_channels=caption_channels,
        model_max_length=model_max_length,
        force_pad_h_for_sp_size=force_pad_h_for_sp_size, qk_norm=True, pred_sigma=False,
        enable_flash_attn=False, enable_layernorm_kernel=False, enable_xformers=False,
        enable_sequence_parallelism=enable_sequence_parallelism, freeze_y_embedder=True,
        with_temp_block=video, use_x_control_embedder=True,
        sequence_parallelism_temporal=False, use_st_cross_attn=False, uncond_cam_in_dim=(3, 7),
        control_depth=control_depth,
        cam_encoder_cls=_EMB + "CamEmbedder",
        cam_encoder_param=dict(input_dim=3, num=7, after_proj=True),
        bbox_embedder_cls=_EMB + "ContinuousBBoxWithTextTempEmbedding",
        bbox_embedder_param=dict(
            n_classes=10, class_token_dim=hidden_size, trainable_class_token=False,
            embedder_num_freq=4, proj_dims=[hidden_size, 512, 512, hidden_size], mode="all-xyz",
            minmax_normalize=False, use_text_encoder_init=True, after_proj=True, sample_id=True,
            **temp_kw),
        map_embedder_cls=_EMB + "MapControlEmbedding",
        map_embedder_param=dict(conditioning_size=[8, map_size, map_size],
                                block_out_channels=[16, 32, 96, 256]),
        map_embedder_downsample_rate=4.5,
        micro_frame_size=None,
        frame_emb_cls=_EMB + "CamEmbedderTemp",
        frame_emb_param=dict(input_dim=3, num=4, after_proj=True, **temp_kw),
        control_skip_cross_view=True, control_skip_temporal=not video,
    )


# named workloads (BASELINE.json configs; App. B of SURVEY.md)
WORKLOADS = {
    # cfg1: tiny, CPU-runnable
    "tiny_1f": dict(model=dict(hidden_size=384, depth=2, num_heads=6, control_depth=1,
                               caption_channels=64, model_max_length=24, map_size=400),
                    b=1, NC=6, t_pix=1, height=224, width=400, n_box=4, n_text=9),
    # stage-1 image configuration (1x224x400 "noTemp" configs): no temporal blocks, un-pooled conditions
    "tiny_1f_image": dict(model=dict(hidden_size=384, depth=2, num_heads=6, control_depth=1, video=False,
                                     caption_channels=64, model_max_length=24, map_size=400),
                          b=1, NC=6, t_pix=1, height=224, width=400, n_box=4, n_text=9),
    "tiny_9f": dict(model=dict(hidden_size=384, depth=2, num_heads=6, control_depth=1,
                               caption_channels=64, model_max_length=24, map_size=400),
                    b=1, NC=6, t_pix=9, height=224, width=400, n_box=4, n_text=9),
    # tiny but with D=72 heads like XL (exercises the padded-head kernels)
    "tiny_d72": dict(model=dict(hidden_size=576, depth=2, num_heads=8, control_depth=1,
                                caption_channels=64, model_max_length=24, map_size=400),
                     b=2, NC=6, t_pix=9, height=224, width=400, n_box=4, n_text=9),
    # long clip at toy width: 65 pixel frames -> T = 17 latent frames (the 8 < T <= 32 temporal kernel at model level)
    "tiny_65f": dict(model=dict(hidden_size=384, depth=2, num_heads=6, control_depth=1,
                                caption_channels=64, model_max_length=24, map_size=400),
                     b=1, NC=6, t_pix=65, height=224, width=400, n_box=4, n_text=9),
    # D = 72 heads with the XL condition length: 40 text + 20 box + cam + frame tokens = 62 keys in cross-attention
    "tiny_l62": dict(model=dict(hidden_size=576, depth=2, num_heads=8, control_depth=1,
                                caption_channels=64, model_max_length=48, map_size=400),
                     b=1, NC=6, t_pix=9, height=224, width=400, n_box=20, n_text=40),
    # cfg2 at b = 1 (one CFG half): the XL-depth golden of the real reference (oracle/make_golden_xl.py)
    "xl_17f_224x400_b1": dict(model=dict(), b=1, NC=6, t_pix=17, height=224, width=400,
                              n_box=20, n_text=40),
    # cfg2: XL/2, 6 views x 17 frames x 224x400, CFG batch-doubled (b=2)
    "xl_17f_224x400": dict(model=dict(), b=2, NC=6, t_pix=17, height=224, width=400,
                           n_box=20, n_text=40),
    # cfg3: XL/2, 6 views x 17 frames x 848x1600
    "xl_17f_848x1600": dict(model=dict(), b=2, NC=6, t_pix=17, height=848, width=1600,
                            n_box=20, n_text=40),
    # cfg4: long video 65 frames 424x800
    "xl_65f_424x800": dict(model=dict(), b=2, NC=6, t_pix=65, height=424, width=800,
                           n_box=20, n_text=40),
}


def latent_frames(t_pix: int) -> int:
    """CogVideoX VAE temporal compression (vae_cogvideox.py:1419-1431): 1->1, 9->3, 17->5, 65->17."""
    return 1 if t_pix == 1 else (t_pix - 1) // 4 + 1


def make_inpaint_inputs(workload: dict, seed: int = 0, kw: dict | None = None):
    """Extra inputs of the BrushNet variant (magicdrive_stdit3.py:1466-1487): the masked pixel video of every view,
    ``x_inpaint [b, 3·NC, Tpix, Hpix, Wpix]``, and its binary mask ``mask_inpaint [b, NC, Tpix, Hpix, Wpix]``."""
    g = torch.Generator().manual_seed(seed + 77)
    b, NC, t_pix = workload["b"], workload["NC"], workload["t_pix"]
    H, W = math.ceil(workload["height"] / 8) * 8, math.ceil(workload["width"] / 8) * 8
    mask = (torch.rand(b, NC, t_pix, H // 8, W // 8, generator=g) > 0.3).float()
    mask = mask.repeat_interleave(8, dim=3).repeat_interleave(8, dim=4)
    x_inpaint = torch.randn(b, 3 * NC, t_pix, H, W, generator=g) * mask.repeat(1, 3, 1, 1, 1)
    return dict(x_inpaint=x_inpaint, mask_inpaint=mask)


def make_sde_inputs(workload: dict, seed: int = 0, kw: dict | None = None):
    """Extra inputs of the SDE-BrushNet variant (magicdrive_stdit3.py:1996-2020): the inpaint branch's own timestep
    ``timestep_inpaint [b]``, the schedule length ``num_timesteps`` and the noise latent mixed into the encoded video,
    ``noise_inpaint_encoded [b, Cin·NC, T, h, w]`` (same packing as ``x``)."""
    kw = kw or model_kwargs(**workload["model"])
    g = torch.Generator().manual_seed(seed + 78)
    b, NC = workload["b"], workload["NC"]
    T = latent_frames(workload["t_pix"])
    h, w = math.ceil(workload["height"] / 8), math.ceil(workload["width"] / 8)
    noise = torch.randn(b, kw["in_channels"] * NC, T, h, w, generator=g)
    return dict(timestep_inpaint=torch.full((b,), 300.0), num_timesteps=1000, noise_inpaint_encoded=noise)


def make_inputs(workload: dict, seed: int = 0, x_mask: str | None = "ones", kw: dict | None = None):
    """Synthetic forward kwargs (CPU tensors, fp32).  ``x_mask``: None | "ones" | "first0"."""
    kw = kw or model_kwargs(**workload["model"])
    g = torch.Generator().manual_seed(seed)
    b, NC, t_pix = workload["b"], workload["NC"], workload["t_pix"]
    H, W = workload["height"], workload["width"]
    C = kw["hidden_size"]
    T = latent_frames(t_pix)
    h, w = math.ceil(H / 8), math.ceil(W / 8)
    L = kw["model_max_length"]
    msz = kw["map_embedder_param"]["conditioning_size"]
    n_box = workload["n_box"]
    B = b * NC

    def rn(*s):
        return torch.randn(*s, generator=g)

    mask = torch.zeros(1, L, dtype=torch.int64)
    mask[:, : workload["n_text"]] = 1
    out = dict(
        x=rn(b, kw["in_channels"] * NC, T, h, w),
        timestep=torch.full((b,), 700.0),
        y=rn(b, 1, L, kw["caption_channels"]),
        maps=(torch.rand(b, t_pix, *msz, generator=g) > 0.5).float(),
        bbox=dict(
            bboxes=rn(B, t_pix, n_box, 8, 3) * 10.0,
            classes=torch.randint(0, 10, (B, t_pix, n_box), generator=g),
            masks=torch.ones(B, t_pix, n_box, dtype=torch.int64),
            box_latent=rn(B, t_pix, n_box, C),
        ),
        cams=rn(B, t_pix, 1, 3, 7),
        rel_pos=rn(B, t_pix, 1, 3, 4),
        fps=torch.full((b,), 12.0),
        height=torch.full((b,), float(H)),
        width=torch.full((b,), float(W)),
        mv_order_map={k: list(v) for k, v in MV_ORDER_MAP.items()} if NC == 6 else
        {i: [(i - 1) % NC, (i + 1) % NC] for i in range(NC)},
        t_order_map=None,
        mask=mask,
        num_frames=t_pix,
    )
    # make part of the boxes "null" (mask 0) so the null-feature path is exercised
    out["bbox"]["masks"][:, :, n_box - 1] = 0
    if x_mask == "ones":
        out["x_mask"] = torch.ones(b, T, dtype=torch.bool)
    elif x_mask == "first0":
        m = torch.ones(b, T, dtype=torch.bool)
        m[:, 0] = False
        out["x_mask"] = m
    else:
        out["x_mask"] = None
    return out


def to_device(obj, device, dtype=None):
    """Moves a (nested) input structure; floating tensors optionally cast."""
    if isinstance(obj, torch.Tensor):
        if dtype is not None and obj.is_floating_point():
            return obj.to(device=device, dtype=dtype)
        return obj.to(device=device)
    if isinstance(obj, dict):
        return {k: to_device(v, device, dtype) for k, v in obj.items()}
    return obj


@torch.no_grad()
def randomize_zero_params(model: torch.nn.Module, seed: int = 1, std: float = 0.02):
    """Re-randomises every all-zero parameter with N(0, std): the reference
    zero-initialises mva_proj / after_proj / before_proj / temporal projections
    (magicdrive_stdit3.py:788-825), which would hide those branches from a parity test."""
    g = torch.Generator().manual_seed(seed)
    n = 0
    for name, p in sorted(model.named_parameters()):
        if p.numel() > 0 and torch.count_nonzero(p) == 0:
            p.copy_(torch.randn(p.shape, generator=g) * std)
            n += 1
    return n


_SMALL_INIT = ("mva_proj.", "after_proj.", "before_proj.", "final_proj.")


@torch.no_grad()
def fill_params(model: torch.nn.Module, seed: int = 0, on_device: bool = False):
    """Deterministic, construction-order-independent weights: every state-dict entry is filled from
    a generator seeded by (seed, crc32(name)), so the reference module, the oracle and the native
    drop-in get bit-identical parameters without shipping a checkpoint.  Scales follow the
    reference init (xavier for matrices) with N(0, 0.02) on biases and on the branches the
    reference zero-initialises (magicdrive_stdit3.py:788-825), so that no branch is silently 0."""
    import zlib
    for name, t in sorted(model.state_dict().items()):
        if not t.is_floating_point() or name.endswith("rope.freqs") or name.endswith("inv_freq"):
            continue
        sd_ = (seed * 1000003 + zlib.crc32(name.encode())) % (2 ** 63)
        if on_device:      # bench.py: same recipe with the device RNG (2 B parameters; values differ from CPU)
            r = torch.randn(t.shape, generator=torch.Generator(device=t.device).manual_seed(sd_), device=t.device)
        else:
            r = torch.randn(t.shape, generator=torch.Generator().manual_seed(sd_))
        small = any(k in name for k in _SMALL_INIT) or (
            "blocks_t." in name and (".attn.proj." in name or ".cross_attn.proj." in name))
        if t.dim() >= 2:
            fan_out, fan_in = t.shape[0], t[0].numel()
            std = 0.02 if small else (2.0 / (fan_in + fan_out)) ** 0.5
            val = r * std
        elif "norm.weight" in name:
            val = 1.0 + 0.1 * r
        else:
            val = r * 0.02
        t.copy_(val.to(t.dtype))
    return model
