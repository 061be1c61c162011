"""Sequence-parallel host logic on CPU: world_size 2 over gloo.  Each rank runs the drop-in with the kernels
replaced by their torch specification (tests/spec_ops.py); the result must equal the single-rank golden vector of
the REAL reference built with ``force_pad_h_for_sp_size`` (the reference's own SP path cannot run on
CPU — `_gather` asserts cuda/npu — so its single-process, H-padded forward is the oracle)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, wl_name, xm, ret):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import spec_ops
    from helpers import build_model, load_golden, rel_l2, use_ops
    from magicdrive_v2_b200 import sp, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    try:
        fix = load_golden(f"{wl_name}_{xm}.pt")
        model, wl, kw = build_model(wl_name, enable_sequence_parallelism=True)
        sp.set_sequence_parallel_group(dist.group.WORLD)
        inp = synth.make_inputs(wl, seed=fix["seed"], x_mask=xm, kw=kw)
        with use_ops(spec_ops):
            out = model(**inp)
        ret[rank] = rel_l2(out, fix["out"])
        # collective helpers keep the reference's contracts (communications.py:60-66,118-155)
        x = torch.arange(2 * 4 * 6 * 3, dtype=torch.float32).view(2, 4, 6, 3) + 1000 * rank
        y = sp.all_to_all(x, dist.group.WORLD, scatter_dim=2, gather_dim=1)
        assert y.shape == (2, 8, 3, 3)
        back = sp.all_to_all(y, dist.group.WORLD, scatter_dim=1, gather_dim=2)
        assert torch.equal(back, x)
        full = sp.gather(sp.split(x, dist.group.WORLD, dim=1), dist.group.WORLD, dim=1)
        ref0 = torch.arange(2 * 4 * 6 * 3, dtype=torch.float32).view(2, 4, 6, 3)
        assert torch.equal(full[:, :2], ref0[:, :2]) and torch.equal(full[:, 2:], (ref0 + 1000)[:, 2:])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("wl_name,xm", [("tiny_9f", "first0"), ("tiny_d72", "first0")])
def test_sequence_parallel_world2_matches_reference_golden(wl_name, xm):
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), wl_name, xm, ret), nprocs=world, join=True)
    assert len(ret) == world
    for r in range(world):
        assert ret[r] < 2e-5, (r, ret[r])


def _worker_brush(rank, world, port, fixture, skip_ca, sde, ret):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import spec_ops
    from helpers import brushnet_inputs, build_brushnet, load_golden, rel_l2, sde_inputs, use_ops
    from magicdrive_v2_b200 import sp
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    try:
        fix = load_golden(fixture)
        model, wl, kw = build_brushnet(fix["workload"], skip_ca, sde=sde, enable_sequence_parallelism=True)
        sp.set_sequence_parallel_group(dist.group.WORLD)
        inp = (sde_inputs if sde else brushnet_inputs)(wl, kw, fix)
        with use_ops(spec_ops):
            out = model(**inp)
        ret[rank] = rel_l2(out, fix["out"])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("fixture,skip_ca,sde", [("brushnet_tiny_9f_None_ca.pt", False, False),
                                                 ("sde_tiny_9f_first0_noca.pt", True, True)])
def test_brushnet_sequence_parallel_world2_matches_reference_golden(fixture, skip_ca, sde):
    """The third (BrushNet / SDE-BrushNet) stream under sequence parallelism: every rank embeds and runs its S-slice of
    the inpaint tokens; result vs the single-process golden of the REAL reference."""
    world = 2
    ret = mp.Manager().dict()
    mp.spawn(_worker_brush, args=(world, _free_port(), fixture, skip_ca, sde, ret), nprocs=world, join=True)
    assert len(ret) == world
    for r in range(world):
        assert ret[r] < 2e-5, (r, ret[r])
