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


def _drop_masks(wl, kw):
    # full-length text mask: the reference asserts it whenever a sample is dropped (magicdrive_stdit3.py:873-874)
    return dict(drop_cond_mask=torch.tensor([1.0, 0.0]),
                drop_frame_mask=torch.tensor([[1.0] * 9, [1.0, 0, 1, 1, 0, 1, 1, 1, 0]]),
                mask=torch.ones(1, kw["model_max_length"], dtype=torch.int64))


# further configurations pinned to the real reference (oracle/make_golden_variants.py -> tests/golden/variant_*.pt):
# name -> (workload, x_mask, constructor overrides, input overrides)
VARIANTS = {
    "ctrl_xview": ("tiny_9f", "first0", dict(control_skip_cross_view=False), None),   # control blocks with cross-view
    "drop_masks": ("tiny_d72", "first0", {}, _drop_masks),        # dropped conditions / frames (null embeddings)
}


def variant_case(name):
    """-> (workload, constructor kwargs, inputs) of a VARIANTS entry."""
    wl_name, xm, kw_over, inp_over = VARIANTS[name]
    wl = synth.WORKLOADS[wl_name]
    kw = synth.model_kwargs(**wl["model"])
    kw.update(kw_over)
    inp = synth.make_inputs(wl, seed=0, x_mask=xm, kw=kw)
    if inp_over is not None:
        inp.update(inp_over(wl, kw))
    return wl_name, kw, inp


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


# ---- sampler fixtures (shared by oracle/make_golden_rflow.py and the rflow tests) ------------------------------
class ToyDenoiser(torch.nn.Module):
    """Stand-in denoiser built from separately rounded fp32 multiplies and adds only, so CPU and GPU agree to the
    bit: enough to pin the sampler's own arithmetic (schedule, CFG, Euler, mask) against the reference."""

    def __init__(self, dtype=torch.bfloat16):
        super().__init__()
        import types
        self.w = torch.nn.Parameter(torch.tensor(0.3, dtype=dtype))
        # what the "slice" samplers read off the denoiser (rf/__init__.py:494-496); eighths, so sums stay exact
        self.camera_embedder = types.SimpleNamespace(uncond_cam=torch.arange(21).view(3, 7).float() / 8 - 1)
        self.frame_embedder = types.SimpleNamespace(uncond_cam=torch.arange(12).view(3, 4).float() / 8 - 0.5)

    @property
    def device(self):
        return self.w.device

    @staticmethod
    def condition_term(cams, rel_pos, bbox, maps):
        """Per-sample scalar that depends on every condition the null builders touch; all inputs are small multiples
        of 1/8, so the sums are exact in fp32 in any order (CPU and GPU agree to the bit)."""
        def tot(t):
            return t.float().flatten(1).sum(dim=1)
        c = tot(cams) + tot(rel_pos[..., :3, :]) + tot(bbox["bboxes"]) + tot(bbox["masks"]) + tot(maps)
        return (c * (1.0 / 64)).view(-1, 1, 1, 1, 1)

    def forward(self, x, timestep, y, x_mask=None, **kwargs):
        xf = x.to(self.w.dtype).float()
        s = y.float().view(-1, 1, 1, 1, 1)
        if "cams" in kwargs:
            s = s + self.condition_term(kwargs["cams"], kwargs["rel_pos"], kwargs["bbox"], kwargs["maps"])
        tt = (timestep.float() * 0.001).view(-1, 1, 1, 1, 1)      # (no division: CUDA divides through a reciprocal)
        out = xf * 0.3
        out = out + s * tt
        if x_mask is not None:
            out = out * x_mask[:, None, :, None, None].float()
            out = out + 0.125
        return out


class ToyBrushDenoiser(ToyDenoiser):
    """Same exact-arithmetic stand-in with the BrushNet call signature (magicdrive_stdit3.py:1466-1487)."""

    def forward(self, x, x_inpaint, mask_inpaint, timestep, y, x_mask=None, **kwargs):
        out = ToyDenoiser.forward(self, x, timestep, y, x_mask=x_mask, **kwargs)
        return out + x_inpaint.float() * mask_inpaint.float() * 0.5


class ToySDEDenoiser(ToyDenoiser):
    """... and with the SDE-BrushNet call signature (magicdrive_stdit3.py:1996-2020)."""

    def forward(self, x, x_inpaint, mask_inpaint, timestep, timestep_inpaint, num_timesteps, y,
                noise_inpaint_encoded=None, x_mask=None, **kwargs):
        out = ToyDenoiser.forward(self, x, timestep, y, x_mask=x_mask, **kwargs)
        out = out + x_inpaint.float() * mask_inpaint.float() * 0.5
        lvl = (timestep_inpaint.float() * (1.0 / 1024)).view(-1, 1, 1, 1, 1) * (num_timesteps * (1.0 / 1024))
        return out + noise_inpaint_encoded.float() * lvl


def rflow_inpaint_inputs(z):
    """(z_inpaint, mask_inpaint) for the toy BrushNet samplers: latent-shaped so the toy can add them elementwise."""
    gen = torch.Generator().manual_seed(13)
    zi = torch.randn(z.shape, generator=gen).to(z.device)
    mi = (torch.rand((z.shape[0], 1) + tuple(z.shape[2:]), generator=gen) > 0.4).float().to(z.device)
    return zi, mi


def rflow_slice_conditions(z):
    """cams / rel_pos / bbox / maps for the "slice" samplers' toy runs (values are eighths, see condition_term)."""
    gen = torch.Generator().manual_seed(17)
    n, T = z.shape[0], z.shape[2]

    def eighths(*shape):
        return (torch.randint(-4, 5, shape, generator=gen).float() / 8).to(z.device)
    return dict(cams=eighths(n, T, 1, 3, 7), rel_pos=eighths(n, T, 1, 4, 4), maps=eighths(n, T, 2, 4, 4),
                bbox=dict(bboxes=eighths(n, T, 2, 8, 3), classes=torch.ones(n, T, 2, dtype=torch.long, device=z.device),
                          masks=torch.ones(n, T, 2, device=z.device)))


# (sampler kind, RFLOW_CASES entry): the un-doubled "slice" family, rf/__init__.py:426-887
RFLOW_SLICE_CASES = [("slice", "plain"), ("slice", "masked"), ("brushnet-slice", "masked"),
                     ("sdebrushnet-slice", "cog_bf16"), ("sdebrushnet-slice", "masked"), ("slice-repaint", "plain"),
                     ("slice-repaint", "cog_bf16")]
REPAINT_IGNORE_T = 0.3


def run_slice_sampler(mod, kind, name, device="cpu"):
    """Runs sampler ``kind`` of module ``mod`` (the reference's ``schedulers.rf`` or ours) on the toy fixtures."""
    kw, z, extra, mask, g = rflow_inputs(name, device=device)
    dt = z.dtype if z.dtype == torch.bfloat16 else torch.float32
    extra = dict(extra, **rflow_slice_conditions(z))
    zi, mi = rflow_inpaint_inputs(z)
    prompts = ["a", "b"][: z.shape[0]]

    class Enc(ToyTextEncoder):
        class t5:                                         # the brushnet slice samplers park the T5 model (:549,558)
            class model:
                device = "cpu"
                to = staticmethod(lambda d: None)

        def encode(self, p):
            return {k: v.to(device) for k, v in super().encode(p).items()}
    common = dict(additional_args=extra, guidance_scale=g, progress=False)
    if kind == "slice":
        return mod.RFLOW_SLICE(**kw).sample(ToyDenoiser(dt).to(device), Enc(), z.clone(), prompts, device, mask=mask,
                                            **common)
    if kind == "brushnet-slice":
        return mod.RFLOW_BRUSHNET_SLICE(**kw).sample(ToyBrushDenoiser(dt).to(device), Enc(), z.clone(), zi, mi,
                                                     prompts, device, mask=mask, **common)
    if kind == "sdebrushnet-slice":
        sched = mod.RFLOW_SDEBRUSHNET_SLICE(inpaint_noise_scale=SDE_NOISE_SCALE, **kw)
        return sched.sample(ToySDEDenoiser(dt).to(device), Enc(), z.clone(), zi, mi, prompts, device, mask=mask,
                            **common)
    assert kind == "slice-repaint"
    gen = torch.Generator().manual_seed(19)
    emask = (torch.rand(z.shape, generator=gen) > 0.5).float().to(device)          # element-wise, shaped like z
    return mod.RFLOW_SLICE_REPAINT(**kw).sample(ToyDenoiser(dt).to(device), Enc(), z.clone(), REPAINT_IGNORE_T, prompts,
                                                device, mask=emask, **common)


RFLOW_BRUSH_CASES = [("brushnet", "plain"), ("brushnet", "masked"), ("sde", "cog_bf16"), ("sde", "masked")]
SDE_NOISE_SCALE = 0.375


class ToyTextEncoder:
    def encode(self, prompts):
        return dict(y=torch.tensor([0.5 + 0.25 * i for i in range(len(prompts))]))

    def null(self, n):
        return torch.full((n,), -0.75)


RFLOW_CASES = {
    # name: (scheduler kwargs, z shape [n, C, T, H, W], z dtype, mask rows or None, guidance)
    "plain": (dict(num_sampling_steps=4, cfg_scale=4.0), (2, 6, 3, 5, 8), torch.float32, None, None),
    "cog_bf16": (dict(num_sampling_steps=5, cfg_scale=2.0, use_timestep_transform=True, cog_style_trans=True),
                 (2, 12, 5, 6, 7), torch.bfloat16, None, None),
    "masked": (dict(num_sampling_steps=6, cfg_scale=2.0, use_timestep_transform=True, cog_style_trans=True),
               (2, 6, 5, 4, 8), torch.bfloat16, [[0.0, 1, 1, 0.5, 1], [1, 1, 0.25, 1, 0]], 3.0),
}


def rflow_inputs(name, device="cpu"):
    kw, shape, dtype, mask, g = RFLOW_CASES[name]
    gen = torch.Generator().manual_seed(7)
    z = torch.randn(shape, generator=gen).to(dtype).to(device)
    n = shape[0]
    extra = dict(height=torch.tensor([848.0] * n, dtype=dtype, device=device),
                 width=torch.tensor([1600.0] * n, dtype=dtype, device=device),
                 num_frames=torch.tensor([17.0] * n, dtype=dtype, device=device))
    mask_t = torch.tensor(mask, device=device) if mask is not None else None
    return kw, z, extra, mask_t, g


class cpu_randn_like:
    """torch.randn_like drawn from a seeded CPU generator wherever the tensor lives (tests only)."""

    def __init__(self, seed=11):
        self.gen = torch.Generator().manual_seed(seed)

    def __call__(self, x):
        return torch.randn(x.shape, generator=self.gen, dtype=x.dtype).to(x.device)


def build_brushnet(workload_name, skip_cross_attn=True, seed=0, sde=False, **extra_kw):
    from magicdrive_v2_b200 import brushnet
    MagicDriveSTDiT3BrushNet = brushnet.MagicDriveSTDiT3SDEBrushNet if sde else brushnet.MagicDriveSTDiT3BrushNet
    wl = synth.WORKLOADS[workload_name]
    kw = synth.model_kwargs(**wl["model"])
    kw["brushnet_skip_cross_attn"] = skip_cross_attn
    kw.update(extra_kw)
    torch.manual_seed(0)
    model = MagicDriveSTDiT3BrushNet(MagicDriveSTDiT3Config(**kw)).eval()
    synth.fill_params(model, seed=seed)
    return model, wl, kw


LORA_KW = dict(use_lora_base_blocks=True, lora_rank=8, lora_alpha=4)      # as oracle/make_golden_brushnet.py


def brushnet_inputs(wl, kw, fix):
    inp = synth.make_inputs(wl, seed=fix["seed"], x_mask=fix["x_mask"], kw=kw)
    inp.update(synth.make_inpaint_inputs(wl, seed=fix["seed"], kw=kw))
    return inp


SDE_CASES = [("sde_tiny_9f_first0_noca.pt", True), ("sde_tiny_9f_None_ca.pt", False)]   # oracle/make_golden_sde.py


def sde_inputs(wl, kw, fix):
    inp = brushnet_inputs(wl, kw, fix)
    inp.update(synth.make_sde_inputs(wl, seed=fix["seed"], kw=kw))
    return inp
