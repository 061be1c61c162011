"""TEST INFRASTRUCTURE — CPU restatement of the rectified-flow sampling loop, eager op by eager op.

Follows ``magicdrivedit/schedulers/rf/__init__.py:38-115`` (``RFLOW.sample``) and
``magicdrivedit/schedulers/rf/rectified_flow.py:42-77`` (``timestep_transform``), ``:157-174`` (``add_noise``).
Pinned: ``tests/test_rflow_cpu.py`` checks it bit-exactly against ``tests/golden/rflow_*.pt``, written by
``oracle/make_golden_rflow.py`` from the REAL reference sampler.  Only ``tests/`` may import this module.
"""
import torch


def timestep_transform(t, height, width, num_frames, num_timesteps, cog_style, base_resolution=512 * 512, scale=1.0):
    # rectified_flow.py:57-77
    t = t / num_timesteps
    ratio_space = ((height * width) / base_resolution).sqrt()
    if num_frames[0] == 1:
        nf = torch.ones_like(num_frames)
    elif cog_style:
        nf = num_frames // 4 + num_frames % 2
    else:
        nf = num_frames // 17 * 5
    ratio = ratio_space * (nf / 1).sqrt() * scale
    return ratio * t / (1 + (ratio - 1) * t) * num_timesteps


def sample(model, model_args, z, num_sampling_steps, num_timesteps=1000, guidance_scale=4.0,
           use_timestep_transform=False, cog_style=False, mask=None, randn_like=torch.randn_like, inpaint=None,
           inpaint_noise_scale=None):
    """model_args already holds the doubled (cond | null) conditions; returns the final latent.
    ``inpaint=(z_inpaint, mask_inpaint)`` selects ``RFLOW_BRUSHNET.sample`` (rf/__init__.py:146-238);
    ``inpaint_noise_scale`` on top of it ``RFLOW_SDEBRUSHNET.sample`` (:291-397)."""
    model_args = dict(model_args)
    n = z.shape[0]
    ts = [torch.tensor([(1.0 - i / num_sampling_steps) * num_timesteps] * n) for i in range(num_sampling_steps)]  # :66-69
    if use_timestep_transform:                                                                                    # :70-74
        ts = [timestep_transform(t, model_args["height"], model_args["width"], model_args["num_frames"],
                                 num_timesteps, cog_style) for t in ts]
    if mask is not None:
        noise_added = mask == 1                                                                                    # :76-78
    for i, t in enumerate(ts):
        if mask is not None:                                                                                       # :83-93
            x0 = z.clone()
            tp = (1 - t.float() / num_timesteps)[:, None, None, None, None]
            x_noise = tp * x0 + (1 - tp) * randn_like(x0)
            upper = mask * num_timesteps >= t.unsqueeze(1)
            model_args["x_mask"] = upper.repeat(2, 1)
            z = torch.where((upper & ~noise_added)[:, None, :, None, None], x_noise, x0)
            noise_added = upper
        if inpaint is None:
            pred = model(torch.cat([z, z], 0), torch.cat([t, t], 0), **model_args)                                 # :95-97
        else:
            zi, mi = (torch.cat([a, a], 0) for a in inpaint)                                                       # :217-218
            if inpaint_noise_scale is None:
                pred = model(torch.cat([z, z], 0), zi, mi, torch.cat([t, t], 0), **model_args)                     # :219
            else:
                t_in = torch.tensor([inpaint_noise_scale * num_timesteps] * n)                                     # :362-363
                nz = randn_like(z)
                pred = model(torch.cat([z, z], 0), zi, mi, torch.cat([t, t], 0), torch.cat([t_in, t_in], 0),
                             num_timesteps, noise_inpaint_encoded=torch.cat([nz, nz], 0), **model_args)            # :372
        pc, pu = pred.chunk(2, dim=0)
        v = pu + guidance_scale * (pc - pu)                                                                        # :101
        dt = (ts[i] - ts[i + 1] if i < len(ts) - 1 else ts[i]) / num_timesteps                                     # :104-105
        z = z + v * dt[:, None, None, None, None]                                                                  # :106
        if mask is not None:
            z = torch.where(upper[:, None, :, None, None], z, x0)                                                  # :108-109
    return z
