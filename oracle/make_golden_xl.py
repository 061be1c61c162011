"""TEST INFRASTRUCTURE — XL-depth and config-shaped goldens of the REAL reference (/root/reference through
oracle/ref_loader.py): tests/golden/{xl_17f_224x400_b1,tiny_65f,tiny_l62}_*.pt.

Run in the build container only:  python oracle/make_golden_xl.py [case ...]
  xl_17f_224x400_b1  MagicDriveSTDiT3-XL/2 (28 + 13 blocks, hidden 1152), 6 views x 17 frames x 224x400, b = 1: one fp32
                     CPU forward of the reference (magicdrive_stdit3.py:1021-1218; minutes on 8 cores) plus its own
                     bf16 forward, so the GPU tests can quote "ours vs fp32" beside "the reference's bf16 vs fp32"
  tiny_65f           65 pixel frames -> T = 17 latent frames (the long-clip temporal kernel at model level)
  tiny_l62           20 boxes + 40 text tokens -> 62 condition tokens, D = 72 heads
Weights come from synth.fill_params (name-keyed seeds), inputs from synth.make_inputs: the fixtures hold outputs only.
"""
import copy
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from magicdrive_v2_b200 import synth  # noqa: E402
from oracle import ref_loader  # noqa: E402

CASES = {"xl_17f_224x400_b1": "first0", "tiny_65f": "first0", "tiny_l62": "ones"}


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    for wl_name in (sys.argv[1:] or CASES):
        xm = CASES[wl_name]
        wl = synth.WORKLOADS[wl_name]
        kw = synth.model_kwargs(**wl["model"])
        torch.manual_seed(0)
        ref = synth.fill_params(ref_loader.build_reference_model(kw).eval(), seed=0)
        inp = synth.make_inputs(wl, seed=0, x_mask=xm, kw=kw)
        t0 = time.time()
        with torch.no_grad():
            out = ref(**inp)
        t_fp32 = time.time() - t0
        fix = dict(workload=wl_name, x_mask=xm, seed=0, out=out.clone(), ref_cpu_fp32_seconds=t_fp32,
                   ref_cpu_threads=torch.get_num_threads())
        name = f"{wl_name}_{xm}.pt"
        torch.save(fix, os.path.join(out_dir, name))             # (the bf16 run below can take long: keep what we have)
        print(name, tuple(out.shape), "mean|v|", out.abs().mean().item(), f"fp32 forward {t_fp32:.1f} s", flush=True)
        t0 = time.time()
        with torch.no_grad():
            ref_bf = copy.deepcopy(ref).to(torch.bfloat16)       # copy: .to() would round rope.freqs in place
            del ref
            out_bf = ref_bf(**synth.to_device(inp, "cpu", torch.bfloat16))
            del ref_bf
        fix["ref_bf16_rel_l2"] = ((out_bf.float() - out).norm() / out.norm()).item()
        fix["ref_cpu_bf16_seconds"] = time.time() - t0
        torch.save(fix, os.path.join(out_dir, name))
        print(name, "ref bf16 gap", fix["ref_bf16_rel_l2"], f"bf16 forward {fix['ref_cpu_bf16_seconds']:.1f} s", flush=True)


if __name__ == "__main__":
    main()
