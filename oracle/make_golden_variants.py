"""TEST INFRASTRUCTURE — writes tests/golden/variant_*.pt: further configurations of the denoiser (tests/helpers.py
VARIANTS: control blocks with cross-view attention, dropped conditions / frames) run through the REAL reference.
Run in the build container only:  python oracle/make_golden_variants.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import helpers  # noqa: E402
from magicdrive_v2_b200 import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402


def main():
    for name in helpers.VARIANTS:
        wl_name, kw, inp = helpers.variant_case(name)
        torch.manual_seed(0)
        ref = synth.fill_params(ref_loader.build_reference_model(kw).eval(), seed=0)
        with torch.no_grad():
            out = ref(**inp)
        torch.save(dict(variant=name, workload=wl_name, seed=0, out=out.clone()),
                   os.path.join(ROOT, "tests", "golden", f"variant_{name}.pt"))
        print(name, tuple(out.shape), "mean|v|", out.abs().mean().item())


if __name__ == "__main__":
    main()
