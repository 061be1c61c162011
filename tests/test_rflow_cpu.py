"""Pins oracle/rflow_oracle.py to the golden latents of the REAL reference sampler, and checks the host half of
the B200 sampler (schedule, registry) against it.  No GPU needed."""
import pytest
import torch

import helpers
from magicdrive_v2_b200 import rflow
from magicdrive_v2_b200.registry import SCHEDULERS
from oracle import rflow_oracle


@pytest.mark.parametrize("name", list(helpers.RFLOW_CASES))
def test_oracle_matches_reference_golden(name):
    kw, z, extra, mask, g = helpers.rflow_inputs(name)
    fix = helpers.load_golden(f"rflow_{name}.pt")
    model = helpers.ToyDenoiser(z.dtype if z.dtype == torch.bfloat16 else torch.float32)
    enc = helpers.ToyTextEncoder()
    n = z.shape[0]
    args = dict(y=torch.cat([enc.encode(["a", "b"][:n])["y"], enc.null(n)], 0), **extra)
    out = rflow_oracle.sample(model, args, z.clone(), kw["num_sampling_steps"],
                              guidance_scale=g if g is not None else kw["cfg_scale"],
                              use_timestep_transform=kw.get("use_timestep_transform", False),
                              cog_style=kw.get("cog_style_trans", False), mask=mask,
                              randn_like=helpers.cpu_randn_like())
    assert out.dtype == fix["out"].dtype and torch.equal(out, fix["out"])       # bit-exact


@pytest.mark.parametrize("name", list(helpers.RFLOW_CASES))
def test_schedule_matches_oracle(name):
    kw, z, extra, mask, g = helpers.rflow_inputs(name)
    sched = SCHEDULERS.build(dict(type="rflow", **kw))
    assert isinstance(sched, rflow.RFLOW)
    ts, dts = sched.schedule(z.shape[0], "cpu", dict(extra))
    steps = kw["num_sampling_steps"]
    ref = [torch.tensor([(1.0 - i / steps) * 1000] * z.shape[0]) for i in range(steps)]
    if kw.get("use_timestep_transform"):
        ref = [rflow_oracle.timestep_transform(t, extra["height"], extra["width"], extra["num_frames"], 1000,
                                               kw.get("cog_style_trans", False)) for t in ref]
    for a, b in zip(ts, ref):
        assert torch.equal(a, b)
    assert dts.shape == (steps, z.shape[0]) and dts.dtype == torch.float32
    for i in range(steps):
        want = (ref[i] - ref[i + 1] if i < steps - 1 else ref[i]) / 1000
        assert torch.equal(dts[i], want.float())
    assert abs(float(dts[:, 0].sum()) - float(ref[0][0]) / 1000) < 1e-6          # the steps cover [t_0, 0]


def test_sampler_has_no_cpu_path():
    kw, z, extra, mask, g = helpers.rflow_inputs("plain")
    sched = rflow.RFLOW(**kw)
    with pytest.raises((RuntimeError, ImportError)):
        sched.sample(helpers.ToyDenoiser(torch.float32), helpers.ToyTextEncoder(), z, ["a", "b"], "cpu",
                     additional_args=dict(extra), progress=False)


def test_add_noise_matches_reference_formula():
    s = rflow.RFlowScheduler(num_timesteps=1000)
    x, nz = torch.randn(2, 3, 4, 5, 6), torch.randn(2, 3, 4, 5, 6)
    t = torch.tensor([250.0, 1000.0])
    tp = (1 - t / 1000)[:, None, None, None, None]
    assert torch.equal(s.add_noise(x, nz, t), tp * x + (1 - tp) * nz)
