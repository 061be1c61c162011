"""TEST INFRASTRUCTURE — writes tests/golden/sde_*.pt by running the REAL reference SDE-BrushNet denoiser
(magicdrive_stdit3.py:1870-2470, imported through oracle/ref_loader.py) on the seeded synthetic workload, with the
inpaint-branch noise supplied by the caller (``noise_inpaint_encoded``; the reference's default draws it with the
un-vendored ``structured_noise`` package, which is a stub here).
Run in the build container only:  python oracle/make_golden_sde.py
"""
import copy
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from magicdrive_v2_b200 import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402

# (workload, x_mask, brushnet_skip_cross_attn)
CASES = [("tiny_9f", "first0", True), ("tiny_9f", None, False)]


def main():
    for wl_name, xm, skip_ca in CASES:
        wl = synth.WORKLOADS[wl_name]
        kw = synth.model_kwargs(**wl["model"])
        kw["brushnet_skip_cross_attn"] = skip_ca
        torch.manual_seed(0)
        ref = synth.fill_params(ref_loader.build_reference_model(kw, brushnet="sde").eval(), seed=0)
        inp = synth.make_inputs(wl, seed=0, x_mask=xm, kw=kw)
        inp.update(synth.make_inpaint_inputs(wl, seed=0, kw=kw))
        inp.update(synth.make_sde_inputs(wl, seed=0, kw=kw))
        with torch.no_grad():
            out = ref(**inp)
            sd = {k: v.clone() for k, v in ref.state_dict().items()}
            out_bf = copy.deepcopy(ref).to(torch.bfloat16)(**synth.to_device(inp, "cpu", torch.bfloat16))
        fix = dict(workload=wl_name, x_mask=xm, seed=0, brushnet_skip_cross_attn=skip_ca, out=out.clone(),
                   ref_bf16_rel_l2=((out_bf - out).norm() / out.norm()).item(),
                   state_shapes={k: tuple(v.shape) for k, v in sd.items()} if skip_ca else None)
        name = f"sde_{wl_name}_{xm}_{'noca' if skip_ca else 'ca'}.pt"
        torch.save(fix, os.path.join(ROOT, "tests", "golden", name))
        print(name, tuple(out.shape), "mean|v|", out.abs().mean().item(), "ref bf16 gap", fix["ref_bf16_rel_l2"])


if __name__ == "__main__":
    main()
