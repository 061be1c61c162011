"""GPU parity of the sampler: mdv2_cfg_euler_step against the eager op sequence (bit-exact) and RFLOW.sample
against the golden latents of the REAL reference sampler (bit-exact: the toy denoiser is exact on both sides)."""
from unittest import mock

import pytest
import torch

import helpers

pytestmark = pytest.mark.gpu


def _eager(pred, z, dt, g, x0=None, keep=None):
    pc, pu = pred.chunk(2, dim=0)
    v = pu + g * (pc - pu)
    out = z + v * dt[:, None, None, None, None]
    if keep is not None:
        out = torch.where(keep[:, None, :, None, None], out, x0)
    return out


@pytest.mark.parametrize("shape", [(2, 96, 5, 106, 200), (1, 6, 3, 5, 7), (3, 4, 2, 3, 4), (2, 16, 1, 28, 50)])
@pytest.mark.parametrize("zdt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("masked", [False, True])
@pytest.mark.parametrize("zin", [None, torch.bfloat16, torch.float32])
def test_cfg_euler_step_bit_exact(shape, zdt, masked, zin):
    from magicdrive_v2_b200 import ops
    gen = torch.Generator(device="cuda").manual_seed(3)
    n = shape[0]
    pred = torch.randn((2 * n,) + shape[1:], device="cuda", generator=gen)
    z = torch.randn(shape, device="cuda", generator=gen).to(zdt)
    dt = torch.rand(n, device="cuda", generator=gen) * 0.1
    x0 = keep = None
    if masked:
        x0 = torch.randn(shape, device="cuda", generator=gen).to(zdt)
        keep = torch.rand((n, shape[2]), device="cuda", generator=gen) > 0.4
    z_in = torch.empty((2 * n,) + shape[1:], device="cuda", dtype=zin) if zin is not None else None
    want = _eager(pred, z, dt, 2.0, x0, keep)
    got = ops.cfg_euler_step(pred, z.clone(), dt, 2.0, x0=x0, keep=keep, z_in=z_in)
    assert got.dtype == torch.float32 and torch.equal(got, want)
    if z_in is not None:
        assert torch.equal(z_in, torch.cat([want, want], 0).to(zin))


def test_cfg_euler_step_empty_and_errors():
    from magicdrive_v2_b200 import ops
    z = torch.zeros((0, 4, 2, 3, 4), device="cuda")
    out = ops.cfg_euler_step(torch.zeros((0, 4, 2, 3, 4), device="cuda"), z, torch.zeros(0, device="cuda"), 1.0)
    assert out.numel() == 0
    with pytest.raises(RuntimeError):
        ops.cfg_euler_step(torch.zeros(2, 1, 1, 2, 2), torch.zeros(1, 1, 1, 2, 2), torch.zeros(1), 1.0)  # CPU tensors


@pytest.mark.parametrize("name", list(helpers.RFLOW_CASES))
def test_sample_matches_reference_golden(name):
    from magicdrive_v2_b200 import ops, rflow
    kw, z, extra, mask, g = helpers.rflow_inputs(name, device="cuda")
    fix = helpers.load_golden(f"rflow_{name}.pt")
    model = helpers.ToyDenoiser(z.dtype if z.dtype == torch.bfloat16 else torch.float32).cuda()

    class Enc(helpers.ToyTextEncoder):
        def encode(self, prompts):
            return {k: v.cuda() for k, v in super().encode(prompts).items()}

    before = ops.launches
    with mock.patch.object(torch, "randn_like", helpers.cpu_randn_like()):
        out = rflow.RFLOW(**kw).sample(model, Enc(), z.clone(), ["a", "b"][: z.shape[0]], "cuda",
                                       additional_args=dict(extra), mask=mask, guidance_scale=g, progress=False)
    assert ops.launches - before == kw["num_sampling_steps"]            # one fused launch per step
    assert out.dtype == torch.float32 and torch.equal(out.cpu(), fix["out"])


@pytest.mark.parametrize("kind,name", helpers.RFLOW_BRUSH_CASES)
def test_brushnet_samplers_match_reference_golden(kind, name):
    """RFLOW_BRUSHNET / RFLOW_SDEBRUSHNET (rf/__init__.py:121-397) against latents of the REAL reference samplers."""
    from magicdrive_v2_b200 import ops, rflow
    kw, z, extra, mask, g = helpers.rflow_inputs(name, device="cuda")
    fix = helpers.load_golden(f"rflow_{kind}_{name}.pt")
    dt = z.dtype if z.dtype == torch.bfloat16 else torch.float32
    zi, mi = helpers.rflow_inpaint_inputs(z)
    if kind == "brushnet":
        sched, model = rflow.RFLOW_BRUSHNET(**kw), helpers.ToyBrushDenoiser(dt).cuda()
    else:
        sched = rflow.RFLOW_SDEBRUSHNET(inpaint_noise_scale=helpers.SDE_NOISE_SCALE, **kw)
        model = helpers.ToySDEDenoiser(dt).cuda()

    class Enc(helpers.ToyTextEncoder):
        def encode(self, prompts):
            return {k: v.cuda() for k, v in super().encode(prompts).items()}

    before = ops.launches
    with mock.patch.object(torch, "randn_like", helpers.cpu_randn_like()):
        out = sched.sample(model, Enc(), z.clone(), zi, mi, ["a", "b"][: z.shape[0]], "cuda",
                           additional_args=dict(extra), mask=mask, guidance_scale=g, progress=False)
    assert ops.launches - before == kw["num_sampling_steps"]
    assert out.dtype == torch.float32 and torch.equal(out.cpu(), fix["out"])


@pytest.mark.parametrize("kind,name", helpers.RFLOW_SLICE_CASES)
def test_slice_samplers_match_reference_golden(kind, name):
    """The un-doubled "slice" family (rf/__init__.py:426-887) against latents of the REAL reference samplers."""
    from magicdrive_v2_b200 import ops, rflow
    fix = helpers.load_golden(f"rflow_{kind}_{name}.pt")
    before = ops.launches
    with mock.patch.object(torch, "randn_like", helpers.cpu_randn_like()):
        out = helpers.run_slice_sampler(rflow, kind, name, device="cuda")
    assert ops.launches - before == helpers.RFLOW_CASES[name][0]["num_sampling_steps"]
    assert out.dtype == torch.float32 and torch.equal(out.cpu(), fix["out"])
