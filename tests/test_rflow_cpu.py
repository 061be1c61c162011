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


@pytest.mark.parametrize("kind,name", helpers.RFLOW_BRUSH_CASES)
def test_brushnet_sampler_oracle_matches_reference_golden(kind, name):
    """RFLOW_BRUSHNET / RFLOW_SDEBRUSHNET restatement vs latents of the REAL reference samplers (bit-exact)."""
    kw, z, extra, mask, g = helpers.rflow_inputs(name)
    fix = helpers.load_golden(f"rflow_{kind}_{name}.pt")
    dt = z.dtype if z.dtype == torch.bfloat16 else torch.float32
    model = helpers.ToyBrushDenoiser(dt) if kind == "brushnet" else helpers.ToySDEDenoiser(dt)
    enc = helpers.ToyTextEncoder()
    n = z.shape[0]
    args = dict(y=torch.cat([enc.encode(["a", "b"][:n])["y"], enc.null(n)], 0), **extra)
    out = rflow_oracle.sample(model, args, z.clone(), kw["num_sampling_steps"],
                              guidance_scale=g if g is not None else kw["cfg_scale"],
                              use_timestep_transform=kw.get("use_timestep_transform", False),
                              cog_style=kw.get("cog_style_trans", False), mask=mask,
                              randn_like=helpers.cpu_randn_like(), inpaint=helpers.rflow_inpaint_inputs(z),
                              inpaint_noise_scale=helpers.SDE_NOISE_SCALE if kind == "sde" else None)
    assert out.dtype == fix["out"].dtype and torch.equal(out, fix["out"])


def test_brushnet_samplers_registered():
    a = SCHEDULERS.build(dict(type="rflow-brushnet", num_sampling_steps=3))
    b = SCHEDULERS.build(dict(type="rflow-sdebrushnet", num_sampling_steps=3, inpaint_noise_scale=0.25))
    assert isinstance(a, rflow.RFLOW_BRUSHNET) and isinstance(b, rflow.RFLOW_SDEBRUSHNET)
    assert b.inpaint_noise_scale == 0.25 and b.num_sampling_steps == 3


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


def _spec_cfg_euler_step(pred, z, dt, g, x0=None, keep=None, z_in=None):
    """Eager specification of mdv2_cfg_euler_step (the op sequence of rf/__init__.py:100-109 + the next cat/cast)."""
    pc, pu = pred.chunk(2, dim=0)
    out = z + (pu + g * (pc - pu)) * dt[:, None, None, None, None]
    if keep is not None:
        out = torch.where(keep.bool()[:, None, :, None, None], out, x0)
    if z_in is not None:
        z_in.copy_(torch.cat([out, out], 0))
    return out.float()


@pytest.mark.parametrize("kind,name", [("base", n) for n in helpers.RFLOW_CASES] + helpers.RFLOW_BRUSH_CASES)
def test_sampler_host_logic_matches_reference_golden(kind, name, monkeypatch):
    """The three samplers' host logic (schedule upload, mask bookkeeping, doubled inputs, RNG draw order, denoiser call
    signatures) with the fused step replaced by its eager specification: bit-exact vs the REAL reference latents."""
    from magicdrive_v2_b200 import ops
    monkeypatch.setattr(ops, "cfg_euler_step", _spec_cfg_euler_step)
    monkeypatch.setattr(torch, "randn_like", helpers.cpu_randn_like())
    kw, z, extra, mask, g = helpers.rflow_inputs(name)
    dt = z.dtype if z.dtype == torch.bfloat16 else torch.float32
    prompts = ["a", "b"][: z.shape[0]]
    common = dict(additional_args=dict(extra), mask=mask, guidance_scale=g, progress=False)
    if kind == "base":
        fix = helpers.load_golden(f"rflow_{name}.pt")
        out = rflow.RFLOW(**kw).sample(helpers.ToyDenoiser(dt), helpers.ToyTextEncoder(), z.clone(), prompts, "cpu",
                                       **common)
    else:
        fix = helpers.load_golden(f"rflow_{kind}_{name}.pt")
        zi, mi = helpers.rflow_inpaint_inputs(z)
        if kind == "brushnet":
            sched, model = rflow.RFLOW_BRUSHNET(**kw), helpers.ToyBrushDenoiser(dt)
        else:
            sched = rflow.RFLOW_SDEBRUSHNET(inpaint_noise_scale=helpers.SDE_NOISE_SCALE, **kw)
            model = helpers.ToySDEDenoiser(dt)
        out = sched.sample(model, helpers.ToyTextEncoder(), z.clone(), zi, mi, prompts, "cpu", **common)
    assert out.dtype == torch.float32 and torch.equal(out, fix["out"])


@pytest.mark.parametrize("kind,name", helpers.RFLOW_SLICE_CASES)
def test_slice_samplers_host_logic_matches_reference_golden(kind, name, monkeypatch):
    """The un-doubled "slice" family (rf/__init__.py:426-887: two denoiser calls per step, null conditions from
    ``replace_with_null_condition``, repaint re-noising) with the fused step replaced by its eager specification:
    bit-exact vs latents of the REAL reference samplers."""
    from magicdrive_v2_b200 import ops
    monkeypatch.setattr(ops, "cfg_euler_step", _spec_cfg_euler_step)
    monkeypatch.setattr(torch, "randn_like", helpers.cpu_randn_like())
    fix = helpers.load_golden(f"rflow_{kind}_{name}.pt")
    out = helpers.run_slice_sampler(rflow, kind, name)
    assert out.dtype == fix["out"].dtype and torch.equal(out, fix["out"])


def test_null_condition_builders():
    """inference_utils.py:144-296 mirrors: structure, the rel_pos row-drop quirk, append / prepend order, errors."""
    from magicdrive_v2_b200.inference_utils import add_null_condition, replace_with_null_condition
    z = torch.zeros(2, 6, 3, 4, 4)
    args = dict(y=torch.ones(2, 1, 5, 8), mask=torch.ones(1, 5), height=torch.tensor([224.0]), fps=12,
                mv_order_map={0: [1]}, **helpers.rflow_slice_conditions(z))
    uc, ur = torch.arange(21.0).view(3, 7), torch.arange(12.0).view(3, 4)
    y_null = torch.zeros(2, 1, 5, 8)
    keys = ["y", "bbox", "cams", "rel_pos", "maps"]
    out = replace_with_null_condition(args, uc, ur, y_null, keys, append=False)
    assert keys == [] and out["y"] is y_null and out["height"] is args["height"] and out["fps"] == 12
    assert out["cams"].shape == args["cams"].shape and torch.equal(out["cams"][1, 2, 0], uc)
    assert out["rel_pos"].shape == (2, 3, 1, 3, 4) and torch.equal(out["rel_pos"][0, 0, 0], ur)
    assert all(torch.count_nonzero(v) == 0 for v in out["bbox"].values()) and torch.count_nonzero(out["maps"]) == 0
    assert torch.equal(out["mask"], args["mask"]) and out["mask"] is not args["mask"]
    both = replace_with_null_condition(args, uc, ur, y_null, ["y", "cams", "maps"], append=True)
    assert both["y"].shape[0] == 4 and torch.equal(both["cams"][:2], args["cams"]) and torch.equal(both["cams"][3, 0, 0], uc)
    assert both["bbox"]["bboxes"].shape[0] == 4 and both["mask"].shape[0] == 2 and both["rel_pos"].shape[0] == 4
    with pytest.raises(RuntimeError):
        replace_with_null_condition(dict(y=args["y"]), uc, ur, y_null, ["y", "cams"])
    dbl = add_null_condition(args, uc, ur, prepend=True, use_map0=True)
    assert torch.equal(dbl["cams"][2:], args["cams"]) and torch.equal(dbl["cams"][0, 0, 0], uc)
    assert dbl["rel_pos"].shape == (4, 3, 1, 3, 4) and torch.count_nonzero(dbl["maps"][:2]) == 0
    assert dbl["y"].shape[0] == 4 and dbl["fps"] == 12
    dbl = add_null_condition(args, uc, ur)
    assert torch.equal(dbl["maps"][:2], dbl["maps"][2:]) and torch.equal(dbl["cams"][:2], args["cams"])


def test_slice_samplers_registered():
    for name, cls in [("rflow-slice", rflow.RFLOW_SLICE), ("rflow-brushnet-slice", rflow.RFLOW_BRUSHNET_SLICE),
                      ("rflow-sdebrushnet-slice", rflow.RFLOW_SDEBRUSHNET_SLICE),
                      ("rflow-slice-repaint", rflow.RFLOW_SLICE_REPAINT)]:
        assert isinstance(SCHEDULERS.build(dict(type=name, num_sampling_steps=2)), cls)
