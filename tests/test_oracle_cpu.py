"""Pins the CPU oracle (oracle/stdit3_oracle.py) to the golden vectors of the REAL reference."""
import pytest
import torch

from helpers import build_model, load_golden, rel_l2
from magicdrive_v2_b200 import synth
from oracle import stdit3_oracle

CASES = [("tiny_1f", "ones"), ("tiny_9f", "ones"), ("tiny_9f", "first0"), ("tiny_9f", None), ("tiny_d72", "first0")]


@pytest.mark.parametrize("wl_name,xm", CASES)
def test_oracle_matches_reference_golden(wl_name, xm):
    fix = load_golden(f"{wl_name}_{xm}.pt")
    model, wl, kw = build_model(wl_name)          # only used as a container of the name-keyed weights
    sd = {k: v.clone() for k, v in model.state_dict().items()}
    inp = synth.make_inputs(wl, seed=fix["seed"], x_mask=xm, kw=kw)
    out = stdit3_oracle.forward(sd, kw, **inp)
    assert out.shape == fix["out"].shape
    assert rel_l2(out, fix["out"]) < 2e-5
