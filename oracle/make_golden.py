"""TEST INFRASTRUCTURE — generates tests/golden/*.pt by running the REAL reference
(/root/reference, imported through oracle/ref_loader.py) on the seeded synthetic workloads.

Run in the build container only:  python oracle/make_golden.py
Weights come from synth.fill_params (name-keyed seeds) and inputs from synth.make_inputs, so the
fixtures hold outputs only.  Also records the bit-exact index maps of SURVEY App. A.8.
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from magicdrive_v2_b200 import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402

CASES = [("tiny_1f", "ones"), ("tiny_1f_image", "ones"), ("tiny_9f", "ones"), ("tiny_9f", "first0"), ("tiny_9f", None),
         ("tiny_d72", "first0")]


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    models = {}
    for wl_name, xm in CASES:
        wl = synth.WORKLOADS[wl_name]
        kw = synth.model_kwargs(**wl["model"])
        if wl_name not in models:
            torch.manual_seed(0)
            m = ref_loader.build_reference_model(kw).eval()
            models[wl_name] = synth.fill_params(m, seed=0)
        ref = models[wl_name]
        inp = synth.make_inputs(wl, seed=0, x_mask=xm, kw=kw)
        with torch.no_grad():
            out = ref(**inp)
            # bf16 run of the reference itself: its own low-precision gap
            import copy
            ref_bf = copy.deepcopy(ref).to(torch.bfloat16)      # copy: .to() would round rope.freqs in place
            out_bf = ref_bf(**synth.to_device(inp, "cpu", torch.bfloat16))
            del ref_bf
        fix = dict(workload=wl_name, x_mask=xm, seed=0, out=out.clone(),
                   ref_bf16_rel_l2=((out_bf - out).norm() / out.norm()).item(),
                   n_state=len(ref.state_dict()),
                   state_keys=sorted(ref.state_dict().keys()) if wl_name == "tiny_9f" and xm == "ones" else None,
                   state_shapes={k: tuple(v.shape) for k, v in ref.state_dict().items()}
                   if wl_name == "tiny_9f" and xm == "ones" else None)
        name = f"{wl_name}_{xm}.pt"
        torch.save(fix, os.path.join(out_dir, name))
        print(name, tuple(out.shape), "mean|v|", out.abs().mean().item(), "ref bf16 gap", fix["ref_bf16_rel_l2"])


if __name__ == "__main__":
    main()
