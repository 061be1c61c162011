"""TEST INFRASTRUCTURE — writes tests/golden/rflow_*.pt by running the REAL reference sampler
(magicdrivedit/schedulers/rf/__init__.py, imported through oracle/ref_loader.py) around the toy denoiser of
tests/helpers.py.  Run in the build container only:  python oracle/make_golden_rflow.py
"""
import os
import sys
from unittest import mock

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from oracle import ref_loader  # noqa: E402


def main():
    rf = ref_loader.load_reference_sampler()
    for name in helpers.RFLOW_CASES:
        kw, z, extra, mask, g = helpers.rflow_inputs(name)
        model = helpers.ToyDenoiser(z.dtype if z.dtype == torch.bfloat16 else torch.float32)
        sched = rf.RFLOW(**kw)
        with mock.patch.object(torch, "randn_like", helpers.cpu_randn_like()):
            out = sched.sample(model, helpers.ToyTextEncoder(), z.clone(), ["a", "b"][: z.shape[0]], "cpu",
                               additional_args=dict(extra), mask=mask, guidance_scale=g, progress=False)
        torch.save(dict(case=name, out=out.clone()), os.path.join(ROOT, "tests", "golden", f"rflow_{name}.pt"))
        print(name, tuple(out.shape), out.dtype, out.float().abs().mean().item())
    for kind, name in helpers.RFLOW_BRUSH_CASES:        # RFLOW_BRUSHNET / RFLOW_SDEBRUSHNET, rf/__init__.py:121-397
        kw, z, extra, mask, g = helpers.rflow_inputs(name)
        dt = z.dtype if z.dtype == torch.bfloat16 else torch.float32
        zi, mi = helpers.rflow_inpaint_inputs(z)
        if kind == "brushnet":
            sched, model = rf.RFLOW_BRUSHNET(**kw), helpers.ToyBrushDenoiser(dt)
        else:
            sched = rf.RFLOW_SDEBRUSHNET(inpaint_noise_scale=helpers.SDE_NOISE_SCALE, **kw)
            model = helpers.ToySDEDenoiser(dt)
        with mock.patch.object(torch, "randn_like", helpers.cpu_randn_like()):
            out = sched.sample(model, helpers.ToyTextEncoder(), z.clone(), zi, mi, ["a", "b"][: z.shape[0]], "cpu",
                               additional_args=dict(extra), mask=mask, guidance_scale=g, progress=False)
        torch.save(dict(case=name, kind=kind, out=out.clone()),
                   os.path.join(ROOT, "tests", "golden", f"rflow_{kind}_{name}.pt"))
        print(kind, name, tuple(out.shape), out.dtype, out.float().abs().mean().item())
    for kind, name in helpers.RFLOW_SLICE_CASES:        # the "slice" family, rf/__init__.py:426-887
        with mock.patch.object(torch, "randn_like", helpers.cpu_randn_like()):
            out = helpers.run_slice_sampler(rf, kind, name)
        torch.save(dict(case=name, kind=kind, out=out.clone()),
                   os.path.join(ROOT, "tests", "golden", f"rflow_{kind}_{name}.pt"))
        print(kind, name, tuple(out.shape), out.dtype, out.float().abs().mean().item())


if __name__ == "__main__":
    main()
