"""Rectified-flow sampler around the denoiser: the caller of the hot path.

Mirror of ``magicdrivedit/schedulers/rf/__init__.py:13-115`` (``RFLOW``) and
``magicdrivedit/schedulers/rf/rectified_flow.py:42-77,157-174`` (``timestep_transform``, ``add_noise``): same
constructor keywords, same ``sample(model, text_encoder, z, prompts, device, ...)`` signature, same arithmetic.
What changes is where the per-step glue runs: the whole schedule (timesteps, transformed timesteps, step sizes) is
computed once and kept on the device, and the tail of every step — CFG combine, Euler update, frame-mask restore and
the doubled, model-dtype input of the NEXT forward — is one launch of ``mdv2_cfg_euler_step`` instead of six eager
passes over the latent (there is no CPU path: the latent must live on a CUDA device).
"""
from __future__ import annotations

import torch

from .registry import SCHEDULERS


def timestep_transform(t, model_kwargs, base_resolution=512 * 512, base_num_frames=1, scale=1.0, num_timesteps=1,
                       cog_style=False):
    """rectified_flow.py:42-77 — resolution / length aware shift of the timestep, same op order and dtypes."""
    for key in ["height", "width", "num_frames"]:
        if model_kwargs[key].dtype == torch.float16:
            model_kwargs[key] = model_kwargs[key].float()
    t = t / num_timesteps
    resolution = model_kwargs["height"] * model_kwargs["width"]
    ratio_space = (resolution / base_resolution).sqrt()
    if model_kwargs["num_frames"][0] == 1:
        num_frames = torch.ones_like(model_kwargs["num_frames"])
    elif cog_style:
        num_frames = model_kwargs["num_frames"] // 4 + model_kwargs["num_frames"] % 2
    else:
        num_frames = model_kwargs["num_frames"] // 17 * 5
    assert (num_frames >= 1).all(), "num_frames cannot be less than 1"
    ratio_time = (num_frames / base_num_frames).sqrt()
    ratio = ratio_space * ratio_time * scale
    assert (ratio > 0).all(), "ratio cannot be 0"
    new_t = ratio * t / (1 + (ratio - 1) * t)
    return new_t * num_timesteps


class RFlowScheduler:
    """The part of rectified_flow.py:80-174 the sampler uses (constructor keywords + ``add_noise``)."""

    def __init__(self, num_timesteps=1000, num_sampling_steps=10, use_discrete_timesteps=False,
                 sample_method="uniform", loc=0.0, scale=1.0, use_timestep_transform=False, transform_scale=1.0,
                 cog_style_trans=False):
        assert sample_method in ["uniform", "logit-normal"]
        assert sample_method == "uniform" or not use_discrete_timesteps, \
            "Only uniform sampling is supported for discrete timesteps"
        self.num_timesteps = num_timesteps
        self.num_sampling_steps = num_sampling_steps
        self.use_discrete_timesteps = use_discrete_timesteps
        self.sample_method = sample_method
        self.use_timestep_transform = use_timestep_transform
        self.transform_scale = transform_scale
        self.cog_style_trans = cog_style_trans

    def add_noise(self, original_samples, noise, timesteps):
        """rectified_flow.py:157-174."""
        timepoints = 1 - timesteps.float() / self.num_timesteps
        timepoints = timepoints[:, None, None, None, None]
        return timepoints * original_samples + (1 - timepoints) * noise

    def training_losses(self, *a, **k):
        raise NotImplementedError("training is outside the denoising hot path (DESIGN.md §7)")


@SCHEDULERS.register_module("rflow")
class RFLOW:
    def __init__(self, num_sampling_steps=10, num_timesteps=1000, cfg_scale=4.0, use_discrete_timesteps=False,
                 use_timestep_transform=False, **kwargs):
        self.num_sampling_steps = num_sampling_steps
        self.num_timesteps = num_timesteps
        self.cfg_scale = cfg_scale
        self.use_discrete_timesteps = use_discrete_timesteps
        self.use_timestep_transform = use_timestep_transform
        self.scheduler = RFlowScheduler(num_timesteps=num_timesteps, num_sampling_steps=num_sampling_steps,
                                        use_discrete_timesteps=use_discrete_timesteps,
                                        use_timestep_transform=use_timestep_transform, **kwargs)

    # -- schedule (rf/__init__.py:66-75,104-105), built once, resident on the device --------------------------
    def schedule(self, n, device, additional_args=None):
        ts = [(1.0 - i / self.num_sampling_steps) * self.num_timesteps for i in range(self.num_sampling_steps)]
        if self.use_discrete_timesteps:
            ts = [int(round(t)) for t in ts]
        ts = [torch.tensor([t] * n, device=device) for t in ts]
        if self.use_timestep_transform:
            ts = [timestep_transform(t, additional_args, num_timesteps=self.num_timesteps,
                                     cog_style=self.scheduler.cog_style_trans) for t in ts]
        dts = [(ts[i] - ts[i + 1] if i < len(ts) - 1 else ts[i]) / self.num_timesteps for i in range(len(ts))]
        return ts, torch.stack([d.float() for d in dts]).contiguous()

    @torch.no_grad()
    def sample(self, model, text_encoder, z, prompts, device, neg_prompts=None, additional_args=None, mask=None,
               guidance_scale=None, progress=True):
        return self._sample(model, text_encoder, z, prompts, device, neg_prompts, additional_args, mask,
                            guidance_scale, progress, lambda z_in, t2, z, args: model(z_in, t2, **args))

    def _device_schedule(self, z, device, additional_args):
        """-> (timesteps, add_noise timepoints [n,1,1,1,1], step sizes [steps, n]) on ``device``.  The schedule is
        computed on the HOST (IEEE division; torch's CUDA kernels divide by a scalar through a rounded reciprocal,
        which would move timesteps by an ulp) and uploaded once."""
        host_args = None
        if additional_args is not None:
            host_args = {k: (v.cpu() if torch.is_tensor(v) else v) for k, v in additional_args.items()
                         if k in ("height", "width", "num_frames")}
        timesteps, dts = self.schedule(z.shape[0], "cpu", host_args)
        keeps = [1 - t.float() / self.num_timesteps for t in timesteps]          # add_noise's timepoints
        return ([t.to(device) for t in timesteps], [k.to(device)[:, None, None, None, None] for k in keeps],
                dts.to(device))

    @staticmethod
    def _progress(steps, progress):
        if progress:
            try:
                from tqdm import tqdm
                return tqdm(steps, leave=False)
            except ImportError:
                pass
        return steps

    def _sample_slice(self, model, text_encoder, z, prompts, device, neg_prompts, additional_args, mask,
                      guidance_scale, progress, call, per_step=None):
        """The "slice" loop (rf/__init__.py:426-523): the conditional and the unconditional half of every step are two
        denoiser calls at the un-doubled batch, for memory-tight runs.  The null conditions do not change between
        steps, so they are built ONCE (the reference rebuilds them every step, ``:493-497``) and the denoiser's
        two-entry condition cache serves both halves from the second step on.  ``call(z, t, args, st)`` with
        ``st = per_step(z)`` drawn once per step (the SDE variant's inpaint noise is shared by the two halves)."""
        from . import ops
        from .inference_utils import replace_with_null_condition
        if guidance_scale is None:
            guidance_scale = self.cfg_scale
        n = len(prompts)
        model_args = text_encoder.encode(prompts)
        if additional_args is not None:
            model_args.update(additional_args)
        if neg_prompts is not None:
            y_null = text_encoder.encode(neg_prompts)["y"]
        else:
            y_null = text_encoder.null(n).to(device)
        null_args = replace_with_null_condition(
            model_args, model.camera_embedder.uncond_cam.to(device), model.frame_embedder.uncond_cam.to(device),
            y_null, ["y", "bbox", "cams", "rel_pos", "maps"], append=False)
        timesteps, keeps, dts = self._device_schedule(z, device, additional_args)
        if mask is not None:
            noise_added = mask == 1
        z = z.contiguous()
        for i in self._progress(range(len(timesteps)), progress):
            t = timesteps[i]
            x0 = keep = None
            cond_i, null_i = model_args, null_args
            if mask is not None:                                    # rf/__init__.py:472-481
                x0 = z.clone()
                x_noise = keeps[i] * x0 + (1 - keeps[i]) * torch.randn_like(x0)
                keep = mask * self.num_timesteps >= t.unsqueeze(1)
                z = torch.where((keep & ~noise_added)[:, None, :, None, None], x_noise, x0).contiguous()
                x0 = x0.to(z.dtype)
                noise_added = keep
                cond_i, null_i = dict(model_args, x_mask=keep), dict(null_args, x_mask=keep)
            halves = []
            st = per_step(z) if per_step is not None else None
            for args in (cond_i, null_i):
                pred = call(z, t, args, st)
                if pred.shape[1] == z.shape[1] * 2:
                    pred = pred.chunk(2, dim=1)[0]
                else:
                    assert pred.shape[1] == z.shape[1]
                halves.append(pred.float())
            z = ops.cfg_euler_step(torch.cat(halves, 0), z, dts[i], guidance_scale, x0=x0, keep=keep)
        return z

    def _sample(self, model, text_encoder, z, prompts, device, neg_prompts, additional_args, mask, guidance_scale,
                progress, call):
        """The sampling loop shared by the three samplers; ``call(z_in, t2, z, model_args)`` is the denoiser call of
        one step (doubled latent in the model dtype, doubled timestep, the fp32 latent, the doubled conditions)."""
        from . import ops                                           # raises without libmdv2.so: no eager fallback
        if guidance_scale is None:
            guidance_scale = self.cfg_scale
        n = len(prompts)
        model_args = text_encoder.encode(prompts)
        if neg_prompts is not None:
            y_null = text_encoder.encode(neg_prompts)["y"]
        else:
            y_null = text_encoder.null(n).to(device)
        model_args["y"] = torch.cat([model_args["y"], y_null], 0)
        if additional_args is not None:
            model_args.update(additional_args)
        timesteps, keeps, dts = self._device_schedule(z, device, additional_args)
        t2 = [torch.cat([t, t], 0) for t in timesteps]
        if mask is not None:
            noise_added = mask == 1

        z = z.contiguous()
        p0 = next(model.parameters(), None) if isinstance(model, torch.nn.Module) else None
        in_dtype = p0.dtype if p0 is not None and p0.dtype in (torch.float32, torch.bfloat16) else torch.float32
        z_in = torch.empty((2 * z.shape[0],) + tuple(z.shape[1:]), device=z.device, dtype=in_dtype)
        z_in[: z.shape[0]].copy_(z)
        z_in[z.shape[0]:].copy_(z)

        for i in self._progress(range(len(timesteps)), progress):
            t = timesteps[i]
            x0 = keep = None
            if mask is not None:                                    # rf/__init__.py:83-93
                mask_t = mask * self.num_timesteps
                x0 = z.clone()
                x_noise = keeps[i] * x0 + (1 - keeps[i]) * torch.randn_like(x0)   # scheduler.add_noise(x0, noise, t)
                keep = mask_t >= t.unsqueeze(1)
                model_args["x_mask"] = keep.repeat(2, 1)
                z = torch.where((keep & ~noise_added)[:, None, :, None, None], x_noise, x0).contiguous()
                x0 = x0.to(z.dtype)
                noise_added = keep
                z_in[: z.shape[0]].copy_(z)
                z_in[z.shape[0]:].copy_(z)
            pred = call(z_in, t2[i], z, model_args)
            if pred.shape[1] == z_in.shape[1] * 2:
                pred = pred.chunk(2, dim=1)[0]
            else:
                assert pred.shape[1] == z_in.shape[1]
            pred = pred.float().contiguous()
            z = ops.cfg_euler_step(pred, z, dts[i], guidance_scale, x0=x0, keep=keep,
                                   z_in=None if mask is not None else z_in)
        return z


def _doubled(t):
    return torch.cat([t, t], 0)


@SCHEDULERS.register_module("rflow-brushnet")
class RFLOW_BRUSHNET(RFLOW):
    """``RFLOW_BRUSHNET`` (rf/__init__.py:121-238): the same loop around the BrushNet denoiser, which also takes the
    masked pixel video and its mask.  The reference concatenates ``[z_inpaint, z_inpaint]`` / ``[mask, mask]`` anew every
    step (``:217-218``); here they are doubled ONCE, so the denoiser's identity-keyed inpaint cache (shallow encoder +
    mask resize) hits on every step after the first."""

    @torch.no_grad()
    def sample(self, model, text_encoder, z, z_inpaint, mask_inpaint, prompts, device, neg_prompts=None,
               additional_args=None, mask=None, guidance_scale=None, progress=True):
        zi, mi = _doubled(z_inpaint), _doubled(mask_inpaint)
        return self._sample(model, text_encoder, z, prompts, device, neg_prompts, additional_args, mask,
                            guidance_scale, progress, lambda z_in, t2, z, args: model(z_in, zi, mi, t2, **args))


@SCHEDULERS.register_module("rflow-sdebrushnet")
class RFLOW_SDEBRUSHNET(RFLOW):
    """``RFLOW_SDEBRUSHNET`` (rf/__init__.py:264-397): the inpaint branch is noised at the fixed level
    ``inpaint_noise_scale`` with a fresh ``randn_like(z)`` every step (``:362-363``)."""

    def __init__(self, num_sampling_steps=10, num_timesteps=1000, cfg_scale=4.0, use_discrete_timesteps=False,
                 use_timestep_transform=False, inpaint_noise_scale=0.0, **kwargs):
        super().__init__(num_sampling_steps, num_timesteps, cfg_scale, use_discrete_timesteps,
                         use_timestep_transform, **kwargs)
        self.inpaint_noise_scale = inpaint_noise_scale

    @torch.no_grad()
    def sample(self, model, text_encoder, z, z_inpaint, mask_inpaint, prompts, device, neg_prompts=None,
               additional_args=None, mask=None, guidance_scale=None, inpaint_noise_scale=None, progress=True):
        if inpaint_noise_scale is None:
            inpaint_noise_scale = self.inpaint_noise_scale
        zi, mi = _doubled(z_inpaint), _doubled(mask_inpaint)
        t_in = torch.tensor([inpaint_noise_scale * self.num_timesteps] * (2 * z.shape[0]), device=device)

        def call(z_in, t2, z, args):
            noise = torch.randn_like(z)
            return model(z_in, zi, mi, t2, t_in, self.num_timesteps, noise_inpaint_encoded=_doubled(noise), **args)
        return self._sample(model, text_encoder, z, prompts, device, neg_prompts, additional_args, mask,
                            guidance_scale, progress, call)


@SCHEDULERS.register_module("rflow-slice")
class RFLOW_SLICE(RFLOW):
    """``RFLOW_SLICE`` (rf/__init__.py:426-523)."""

    @torch.no_grad()
    def sample(self, model, text_encoder, z, prompts, device, neg_prompts=None, additional_args=None, mask=None,
               guidance_scale=None, progress=True):
        return self._sample_slice(model, text_encoder, z, prompts, device, neg_prompts, additional_args, mask,
                                  guidance_scale, progress, lambda z, t, args, st: model(z, t, **args))


@SCHEDULERS.register_module("rflow-brushnet-slice")
class RFLOW_BRUSHNET_SLICE(RFLOW_BRUSHNET):
    """``RFLOW_BRUSHNET_SLICE`` (rf/__init__.py:525-643); the text-encoder / model device shuffling around the loop
    (``:549,558-559,642-643``) is an offloading aid of the reference's pipeline and not reproduced."""

    @torch.no_grad()
    def sample(self, model, text_encoder, z, z_inpaint, mask_inpaint, prompts, device, neg_prompts=None,
               additional_args=None, mask=None, guidance_scale=None, progress=True):
        return self._sample_slice(model, text_encoder, z, prompts, device, neg_prompts, additional_args, mask,
                                  guidance_scale, progress,
                                  lambda z, t, args, st: model(z, z_inpaint, mask_inpaint, t, **args))


@SCHEDULERS.register_module("rflow-sdebrushnet-slice")
class RFLOW_SDEBRUSHNET_SLICE(RFLOW_SDEBRUSHNET):
    """``RFLOW_SDEBRUSHNET_SLICE`` (rf/__init__.py:646-777): one ``randn_like(z)`` per step, shared by both halves."""

    @torch.no_grad()
    def sample(self, model, text_encoder, z, z_inpaint, mask_inpaint, prompts, device, neg_prompts=None,
               additional_args=None, mask=None, guidance_scale=None, inpaint_noise_scale=None, progress=True):
        if inpaint_noise_scale is None:
            inpaint_noise_scale = self.inpaint_noise_scale
        t_in = torch.tensor([inpaint_noise_scale * self.num_timesteps] * z.shape[0], device=device)

        def call(z, t, args, noise):
            return model(z, z_inpaint, mask_inpaint, t, t_in, self.num_timesteps, noise_inpaint_encoded=noise, **args)
        return self._sample_slice(model, text_encoder, z, prompts, device, neg_prompts, additional_args, mask,
                                  guidance_scale, progress, call, per_step=torch.randn_like)


@SCHEDULERS.register_module("rflow-slice-repaint")
class RFLOW_SLICE_REPAINT(RFLOW):
    """``RFLOW_SLICE_REPAINT`` (rf/__init__.py:780-887): sampling starts from fresh noise; after every step whose
    timestep is at or above ``ignore_mask_timestep`` the masked region is re-imposed from the reference latent noised
    to the NEXT timestep (``z = noised(ref)·mask + z·(1 - mask)``, element-wise ``mask`` shaped like ``z``)."""

    @torch.no_grad()
    def sample(self, model, text_encoder, z, ignore_mask_timestep, prompts, device, neg_prompts=None,
               additional_args=None, mask=None, guidance_scale=None, progress=True):
        from . import ops
        from .inference_utils import replace_with_null_condition
        if guidance_scale is None:
            guidance_scale = self.cfg_scale
        n = len(prompts)
        model_args = text_encoder.encode(prompts)
        if additional_args is not None:
            model_args.update(additional_args)
        if neg_prompts is not None:
            y_null = text_encoder.encode(neg_prompts)["y"]
        else:
            y_null = text_encoder.null(n).to(device)
        null_args = replace_with_null_condition(
            model_args, model.camera_embedder.uncond_cam.to(device), model.frame_embedder.uncond_cam.to(device),
            y_null, ["y", "bbox", "cams", "rel_pos", "maps"], append=False)
        timesteps, keeps, dts = self._device_schedule(z, device, additional_args)
        t_first = [t[0].cpu() for t in timesteps]                   # host copies: no per-step device read-back
        keeps.append(torch.ones(1, device=device)[:, None, None, None, None])    # "next t" after the last step is 0
        if mask is not None:
            assert mask.shape == z.shape
            ref_z = z.clone()
            z = torch.randn_like(ref_z)
        z = z.contiguous()
        all_frames = torch.ones([1, z.shape[2]], dtype=torch.bool, device=device)
        cond_args, null_args = dict(model_args, x_mask=all_frames), dict(null_args, x_mask=all_frames)
        for i in self._progress(range(len(timesteps)), progress):
            halves = []
            for args in (cond_args, null_args):
                pred = model(z, timesteps[i], **args)
                if pred.shape[1] == z.shape[1] * 2:
                    pred = pred.chunk(2, dim=1)[0]
                else:
                    assert pred.shape[1] == z.shape[1]
                halves.append(pred.float())
            z = ops.cfg_euler_step(torch.cat(halves, 0), z, dts[i], guidance_scale)
            if mask is not None and bool(t_first[i] >= ignore_mask_timestep * self.num_timesteps):
                x0 = ref_z.clone()
                x_noise = keeps[i + 1] * x0 + (1 - keeps[i + 1]) * torch.randn_like(x0)
                z = (x_noise * mask + z * (1 - mask)).contiguous()
        return z
