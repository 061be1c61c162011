"""Pins the CPU oracle (oracle/stdit3_oracle.py) to the golden vectors of the REAL reference."""
import pytest
import torch

from helpers import build_model, load_golden, rel_l2
from magicdrive_v2_b200 import synth
from oracle import stdit3_oracle

CASES = [("tiny_1f", "ones"), ("tiny_1f_image", "ones"), ("tiny_9f", "ones"), ("tiny_9f", "first0"), ("tiny_9f", None), ("tiny_d72", "first0"),
         ("tiny_65f", "first0"), ("tiny_l62", "ones")]


@pytest.mark.parametrize("wl_name,xm", CASES)
def test_oracle_matches_reference_golden(wl_name, xm):
    fix = load_golden(f"{wl_name}_{xm}.pt")
    model, wl, kw = build_model(wl_name)          # only used as a container of the name-keyed weights
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    inp = synth.make_inputs(wl, seed=fix["seed"], x_mask=xm, kw=kw)
    out = stdit3_oracle.forward(sd, kw, **inp)
    assert out.shape == fix["out"].shape
    assert rel_l2(out, fix["out"]) < 2e-5


BRUSH_CASES = [("tiny_9f", "first0", True, False), ("tiny_9f", None, False, False), ("tiny_9f", "ones", True, True)]


def _brush(wl_name, skip_ca, lora=False):
    from magicdrive_v2_b200.brushnet import MagicDriveSTDiT3BrushNet
    from magicdrive_v2_b200.config import MagicDriveSTDiT3Config
    wl = synth.WORKLOADS[wl_name]
    kw = synth.model_kwargs(**wl["model"])
    kw["brushnet_skip_cross_attn"] = skip_ca
    if lora:
        kw.update(use_lora_base_blocks=True, lora_rank=8, lora_alpha=4)
    torch.manual_seed(0)
    model = synth.fill_params(MagicDriveSTDiT3BrushNet(MagicDriveSTDiT3Config(**kw)).eval(), seed=0)
    return model, wl, kw


@pytest.mark.parametrize("wl_name,xm,skip_ca,lora", BRUSH_CASES)
def test_brushnet_oracle_matches_reference_golden(wl_name, xm, skip_ca, lora):
    """BrushNet variant (magicdrive_stdit3.py:1314-1867, with and without LoRA adapters on the base blocks): oracle
    vs the output of the REAL reference."""
    fix = load_golden(f"brushnet_{wl_name}_{xm}_{'noca' if skip_ca else 'ca'}{'_lora' if lora else ''}.pt")
    model, wl, kw = _brush(wl_name, skip_ca, lora)
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    if fix["state_shapes"] is not None:              # same key set and shapes as the reference module
        assert {k: tuple(v.shape) for k, v in sd.items()} == fix["state_shapes"]
    inp = synth.make_inputs(wl, seed=fix["seed"], x_mask=xm, kw=kw)
    inp.update(synth.make_inpaint_inputs(wl, seed=fix["seed"], kw=kw))
    out = stdit3_oracle.forward(sd, kw, **inp)
    assert out.shape == fix["out"].shape
    assert rel_l2(out, fix["out"]) < 2e-5


@pytest.mark.parametrize("fixture,skip_ca", [("sde_tiny_9f_first0_noca.pt", True), ("sde_tiny_9f_None_ca.pt", False)])
def test_sde_brushnet_oracle_matches_reference_golden(fixture, skip_ca):
    """SDE-BrushNet variant (magicdrive_stdit3.py:1870-2470: noised inpaint encoding at its own timestep, brushnet
    blocks modulated by t_combine_block): oracle vs the output of the REAL reference, same state-dict layout."""
    from helpers import build_brushnet, sde_inputs
    fix = load_golden(fixture)
    model, wl, kw = build_brushnet(fix["workload"], skip_ca, sde=True)
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    if fix["state_shapes"] is not None:
        assert {k: tuple(v.shape) for k, v in sd.items()} == fix["state_shapes"]
    out = stdit3_oracle.forward(sd, kw, **sde_inputs(wl, kw, fix))
    assert out.shape == fix["out"].shape
    assert rel_l2(out, fix["out"]) < 2e-5
