"""GPU parity of the whole drop-in (through libmdv2) against the golden vectors produced by the REAL
reference (tests/golden, oracle/make_golden.py) — same name-keyed weights, same seeded inputs.
Tolerances (BASELINE.json north_star): relative L2 <= 1e-4 in fp32, <= 1e-2 in bf16 (the reference's
own bf16-vs-fp32 gap on these cases is ~1.0e-2 and is stored in each fixture)."""
import pytest
import torch

from helpers import build_model, load_golden, rel_l2
from magicdrive_v2_b200 import synth

pytestmark = pytest.mark.gpu
CASES = [("tiny_1f", "ones"), ("tiny_1f_image", "ones"), ("tiny_9f", "ones"), ("tiny_9f", "first0"), ("tiny_9f", None),
         ("tiny_d72", "first0"),
         ("tiny_65f", "first0"),     # 65 pixel frames -> T = 17: the long-clip temporal kernel at model level
         ("tiny_l62", "ones")]       # 20 boxes + 40 text tokens -> 62 condition tokens, D = 72 heads


def _run(wl_name, xm, dtype):
    fix = load_golden(f"{wl_name}_{xm}.pt")
    model, wl, kw = build_model(wl_name)
    model = model.to("cuda", dtype)
    inp = synth.to_device(synth.make_inputs(wl, seed=fix["seed"], x_mask=xm, kw=kw), "cuda")
    out = model(**inp)
    torch.cuda.synchronize()
    assert out.dtype == torch.float32 and out.shape == fix["out"].shape
    assert torch.isfinite(out).all()
    return rel_l2(out.cpu(), fix["out"]), fix


@pytest.mark.parametrize("wl_name,xm", CASES)
def test_fp32_matches_reference(wl_name, xm):
    err, _ = _run(wl_name, xm, torch.float32)
    print(f"fp32 {wl_name}/{xm}: rel L2 {err:.3e}")
    assert err < 1e-4


@pytest.mark.parametrize("wl_name,xm", CASES)
def test_bf16_matches_reference(wl_name, xm):
    err, fix = _run(wl_name, xm, torch.bfloat16)
    print(f"bf16 {wl_name}/{xm}: rel L2 {err:.3e} (reference's own bf16 gap {fix['ref_bf16_rel_l2']:.3e})")
    # north star: <= 1e-2.  Where the REFERENCE's own bf16 forward is already further than that from its fp32 forward
    # (tiny_l62: 1.06e-2 — the bound is of the order of bf16 noise), no bf16 implementation can be asked for more than
    # "no further from the fp32 reference than the reference's own bf16 run": that is then the bound, not a wider one.
    ref_gap = fix["ref_bf16_rel_l2"]
    assert err < (1e-2 if ref_gap <= 1e-2 else ref_gap)


def test_native_library_is_the_path_that_runs():
    from magicdrive_v2_b200 import ops
    model, wl, kw = build_model("tiny_1f")
    model = model.to("cuda", torch.bfloat16)
    inp = synth.to_device(synth.make_inputs(wl, kw=kw), "cuda")
    before = ops.launches
    model(**inp)
    assert ops.launches - before > 50            # every hot-path step is a libmdv2 launch
    assert type(model).__mro__[0].__name__ == "MagicDriveSTDiT3" and type(model)._ops is ops



@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 1e-2)])
def test_condition_cache_on_the_real_model(dtype, tol):
    """SURVEY §8 f1 on the GPU: with the condition / cross-attention K-V cache on, the second denoising step of a
    sample (same condition tensors, a new latent and timestep) re-uses the encodings — fewer libmdv2 launches, the
    cache entry untouched — and still matches the reference golden; changed conditions miss."""
    from magicdrive_v2_b200 import ops
    fix = load_golden("tiny_9f_first0.pt")
    model, wl, kw = build_model("tiny_9f")
    model = model.to("cuda", dtype)
    model.cache_conditions = True
    inp = synth.to_device(synth.make_inputs(wl, seed=fix["seed"], x_mask="first0", kw=kw), "cuda")
    n0 = ops.launches
    out1 = model(**inp)
    n1 = ops.launches
    entry = model._cond_cache[0][1]
    other = dict(inp, x=inp["x"] * 0.5, timestep=inp["timestep"] * 0.25)     # the next step of the same sample
    model(**other)
    n2 = ops.launches
    assert model._cond_cache[0][1] is entry                      # hit: nothing was re-encoded
    assert n2 - n1 < n1 - n0                                     # ... and the kv_linear / map GEMMs did not run again
    out2 = model(**inp)                                          # the first step again, now from the cache
    torch.cuda.synchronize()
    assert torch.equal(out1, out2)
    assert rel_l2(out2.cpu(), fix["out"]) < tol
    changed = dict(inp, y=inp["y"].clone())                      # another condition tensor: a miss, a new entry
    out3 = model(**changed)
    assert model._cond_cache[0][1] is not entry
    assert torch.equal(out3, out1)
    model.cache_conditions = False
    out4 = model(**inp)
    assert torch.equal(out4, out1)


def test_captured_step_replays_the_eager_step():
    """one denoising step as a CUDA graph — replays must be bit-identical to the eager forward for every
    new latent / timestep, with the conditions encoded once."""
    model, wl, kw = build_model("tiny_9f")
    model = model.to("cuda", torch.bfloat16)
    model.cache_conditions = True
    inp = synth.to_device(synth.make_inputs(wl, seed=3, x_mask="first0", kw=kw), "cuda")
    step = model.capture_step(**inp)
    for i, (sx, st) in enumerate(((1.0, 1.0), (0.5, 0.25), (-2.0, 0.75))):
        x, t = inp["x"] * sx, inp["timestep"] * st
        want = model(**dict(inp, x=x, timestep=t)).clone()
        got = step(x, t)
        torch.cuda.synchronize()
        assert torch.isfinite(got).all() and torch.equal(got, want), i

BRUSH_CASES = [("brushnet_tiny_9f_first0_noca.pt", True), ("brushnet_tiny_9f_None_ca.pt", False),
               ("brushnet_tiny_9f_ones_noca_lora.pt", True)]


@pytest.mark.parametrize("fixture,skip_ca", BRUSH_CASES)
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 1e-2)])
def test_brushnet_matches_reference(fixture, skip_ca, dtype, tol):
    """BrushNet variant (third stream of control-style blocks, magicdrive_stdit3.py:1314-1867) through libmdv2
    against the output of the REAL reference module."""
    from helpers import LORA_KW, brushnet_inputs, build_brushnet
    fix = load_golden(fixture)
    model, wl, kw = build_brushnet(fix["workload"], skip_ca, **(LORA_KW if fix.get("lora") else {}))
    model = model.to("cuda", dtype)
    inp = synth.to_device(brushnet_inputs(wl, kw, fix), "cuda")
    out = model(**inp)
    torch.cuda.synchronize()
    assert out.dtype == torch.float32 and out.shape == fix["out"].shape and torch.isfinite(out).all()
    err = rel_l2(out.cpu(), fix["out"])
    print(f"brushnet {fixture} {dtype}: rel L2 {err:.3e} (reference's own bf16 gap {fix['ref_bf16_rel_l2']:.3e})")
    assert err < tol


@pytest.mark.parametrize("fixture,skip_ca", [("sde_tiny_9f_first0_noca.pt", True), ("sde_tiny_9f_None_ca.pt", False)])
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 1e-2)])
def test_sde_brushnet_matches_reference(fixture, skip_ca, dtype, tol):
    """SDE-BrushNet variant (magicdrive_stdit3.py:1870-2470) through libmdv2 against the REAL reference module."""
    from helpers import build_brushnet, sde_inputs
    fix = load_golden(fixture)
    model, wl, kw = build_brushnet(fix["workload"], skip_ca, sde=True)
    model = model.to("cuda", dtype)
    inp = synth.to_device(sde_inputs(wl, kw, fix), "cuda")
    out = model(**inp)
    torch.cuda.synchronize()
    assert out.dtype == torch.float32 and out.shape == fix["out"].shape and torch.isfinite(out).all()
    err = rel_l2(out.cpu(), fix["out"])
    print(f"sde {fixture} {dtype}: rel L2 {err:.3e} (reference's own bf16 gap {fix['ref_bf16_rel_l2']:.3e})")
    assert err < tol


@pytest.mark.parametrize("name", ["ctrl_xview", "drop_masks"])
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 1e-2)])
def test_variants_match_reference(name, dtype, tol):
    """Control blocks with cross-view attention; dropped conditions / frames — through libmdv2 against the REAL
    reference, at the north-star tolerances."""
    from helpers import variant_case
    wl_name, kw, inp = variant_case(name)
    fix = load_golden(f"variant_{name}.pt")
    model, _, _ = build_model(wl_name, control_skip_cross_view=kw["control_skip_cross_view"])
    model = model.to("cuda", dtype)
    out = model(**synth.to_device(inp, "cuda"))
    torch.cuda.synchronize()
    err = rel_l2(out.cpu(), fix["out"])
    print(f"variant {name} {dtype}: rel L2 {err:.3e}")
    assert torch.isfinite(out).all() and err < tol


def _xl_case(dtype):
    fix = load_golden("xl_17f_224x400_b1_first0.pt")
    model, wl, kw = build_model("xl_17f_224x400_b1")          # 28 + 13 blocks, hidden 1152; CPU-seeded weights
    model = model.to("cuda", dtype)
    inp = synth.to_device(synth.make_inputs(wl, seed=fix["seed"], x_mask="first0", kw=kw), "cuda")
    out = model(**inp)
    torch.cuda.synchronize()
    assert out.dtype == torch.float32 and out.shape == fix["out"].shape and torch.isfinite(out).all()
    return rel_l2(out.cpu(), fix["out"]), fix


def test_xl_depth_fp32_matches_reference():
    """MagicDriveSTDiT3-XL/2 at full depth (configs[1] at b = 1: 6 views x 17 frames x 224x400) against ONE fp32 CPU
    forward of the REAL reference (oracle/make_golden_xl.py; magicdrive_stdit3.py:1021-1218)."""
    err, _ = _xl_case(torch.float32)
    print(f"XL fp32: rel L2 {err:.3e}")
    assert err < 1e-4


def test_xl_depth_bf16_vs_reference():
    """The same in bf16.  At depth 28 the REFERENCE's own bf16 forward is 2.36e-2 away from its fp32 forward (stored
    in the fixture), so 1e-2 is not a bound any bf16 implementation of this depth meets; what is asserted is that the
    bf16 drop-in is no further from the fp32 reference than the reference's own bf16 run, and the 1e-2 figure of the
    north star is reported beside it (and asserted on every depth-2 configuration above)."""
    err, fix = _xl_case(torch.bfloat16)
    print(f"XL bf16: rel L2 {err:.3e} vs fp32 reference; the reference's own bf16 forward: {fix['ref_bf16_rel_l2']:.3e}")
    assert err <= fix["ref_bf16_rel_l2"]
