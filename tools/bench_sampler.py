"""Secondary measurement (BASELINE.json configs[4] shape, on ONE GPU): the full rectified-flow sampling loop —
``rflow-brushnet`` sampler around the XL/2 BrushNet denoiser at 6 views x 17 frames x 424x800, CFG batch-doubled,
random-init weights, synthetic conditions — timed on the device with CUDA events around ``sample()``.  Not the bench
line (bench.py is); prints one JSON line.   python tools/bench_sampler.py [--sampling-steps 30] [--sde]
"""
import argparse
import json
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class SyntheticTextEncoder:
    """Stands in for the T5 wrapper (``t5.py:195-197,303-330``): fixed embeddings, same ``encode`` / ``null`` contract."""

    def __init__(self, model, n_text, device, dtype):
        self.model, self.n_text, self.device, self.dtype = model, n_text, device, dtype

    def encode(self, prompts):
        L, Cc = self.model.y_embedder.y_embedding.shape
        g = torch.Generator().manual_seed(5)
        mask = torch.zeros(len(prompts), L, dtype=torch.int64)
        mask[:, : self.n_text] = 1
        return dict(y=torch.randn(len(prompts), 1, L, Cc, generator=g).to(self.device, self.dtype),
                    mask=mask.to(self.device))

    def null(self, n):
        return self.model.y_embedder.y_embedding[None].repeat(n, 1, 1)[:, None]


def build(args, dev="cuda:0", dtype=torch.bfloat16, on_device=True, sp=False):
    """-> (sampler, denoiser, text encoder, z, inpaint inputs, doubled conditions, workload) the way
    scripts/inference_magicdrive.py assembles them before ``scheduler.sample``."""
    from magicdrive_v2_b200 import brushnet, rflow, synth
    from magicdrive_v2_b200.config import MagicDriveSTDiT3Config
    from magicdrive_v2_b200.inference_utils import add_null_condition
    wl = dict(synth.WORKLOADS["tiny_9f" if args.tiny else "xl_17f_224x400"], b=1, t_pix=args.frames,
              height=args.height, width=args.width)
    kw = synth.model_kwargs(**dict(wl["model"], force_pad_h_for_sp_size=getattr(args, "pad_sp", 4),
                                   enable_sequence_parallelism=sp))
    cls = brushnet.MagicDriveSTDiT3SDEBrushNet if args.sde else brushnet.MagicDriveSTDiT3BrushNet
    torch.manual_seed(0)
    with torch.device(dev):
        model = cls(MagicDriveSTDiT3Config(**kw)).eval()
    synth.fill_params(model, seed=0, on_device=on_device)
    model = model.to(device=dev, dtype=dtype)
    model.cache_conditions = True                      # conditions / inpaint encodings are computed once per sample

    inp = synth.to_device(synth.make_inputs(wl, seed=0, x_mask=None, kw=kw), dev)
    pix = synth.to_device(synth.make_inpaint_inputs(wl, seed=0, kw=kw), dev, dtype)
    z = torch.randn(inp["x"].shape, generator=torch.Generator().manual_seed(3)).to(dev)
    cond = dict(maps=inp["maps"].to(dtype), bbox={k: (v.to(dtype) if v.is_floating_point() else v)
                                                  for k, v in inp["bbox"].items()},
                cams=inp["cams"].to(dtype),
                rel_pos=torch.nn.functional.pad(inp["rel_pos"], (0, 0, 0, 1)).to(dtype),      # 4x4 poses, as loaded
                fps=inp["fps"], height=inp["height"], width=inp["width"],
                num_frames=torch.full((1,), float(args.frames), device=dev), mv_order_map=inp["mv_order_map"],
                t_order_map=None)
    # what scripts/inference_magicdrive.py does before calling the sampler: null conditions appended (CFG half 2)
    cond = add_null_condition(cond, model.camera_embedder.uncond_cam.to(dev), model.frame_embedder.uncond_cam.to(dev),
                              prepend=False, use_map0=True)
    enc = SyntheticTextEncoder(model, wl["n_text"], dev, dtype)
    skw = dict(num_sampling_steps=args.sampling_steps, cfg_scale=2.0, use_timestep_transform=True, cog_style_trans=True)
    sched = rflow.RFLOW_SDEBRUSHNET(inpaint_noise_scale=0.1, **skw) if args.sde else rflow.RFLOW_BRUSHNET(**skw)
    return sched, model, enc, z, pix, cond, wl


def parse(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--sampling-steps", type=int, default=30)
    ap.add_argument("--height", type=int, default=424)
    ap.add_argument("--width", type=int, default=800)
    ap.add_argument("--frames", type=int, default=17)
    ap.add_argument("--sde", action="store_true")
    ap.add_argument("--tiny", action="store_true", help="depth-2 / hidden-384 denoiser (plumbing check)")
    ap.add_argument("--pad-sp", dest="pad_sp", type=int, default=4, help="force_pad_h_for_sp_size of the denoiser")
    return ap.parse_args(argv)


def main():
    """One process per GPU under ``python -m torch.distributed.run`` (sequence parallelism over the ranks, as
    scripts/inference_magicdrive.py runs the reference on 8 GPUs); a single process otherwise."""
    args = parse()
    from magicdrive_v2_b200 import ops, synth
    world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", 1), ("RANK", 0), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    dev = f"cuda:{local}"
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(dev))
    sched, model, enc, z, pix, cond, wl = build(args, dev, sp=world > 1)
    if world > 1:
        model.sp_group = dist.group.WORLD
    cls = type(model)

    def run(steps):
        sched.num_sampling_steps = steps
        return sched.sample(model, enc, z.clone(), pix["x_inpaint"], pix["mask_inpaint"], ["a driving scene"], dev,
                            additional_args=dict(cond), progress=False)
    run(2)                                             # warm-up: workspaces, condition + inpaint caches
    torch.cuda.synchronize()
    if dist:
        dist.barrier()
    n0 = ops.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    out = run(args.sampling_steps)
    e1.record()
    torch.cuda.synchronize()
    t_ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if dist:
        dist.all_reduce(t_ms, op=dist.ReduceOp.MAX)    # device time, max over the ranks
    ms = t_ms.item()
    assert torch.isfinite(out).all()
    T = synth.latent_frames(args.frames)
    if rank == 0:
        from magicdrive_v2_b200 import sp as _sp
        print(json.dumps(dict(
            metric="rectified-flow sampling loop, denoise steps/sec (device-timed)", value=args.sampling_steps / (ms * 1e-3),
            unit="steps/s", n_gpus=world, sampling_steps=args.sampling_steps, ms_per_step=ms / args.sampling_steps,
            ms_total=ms, dtype="bf16", data="synthetic", gpu_launches=ops.launches - n0,
            config=dict(workload=f"{'sde-' if args.sde else ''}brushnet_{args.frames}f_{args.height}x{args.width}",
                        denoiser=("tiny " if args.tiny else "XL/2 ") + cls.__name__, sampler=type(sched).__name__,
                        views=6, latent_frames=T, cfg_batch=2, parallelism=f"sp{world}" if world > 1 else "single",
                        sp_exchange=_sp.exchange_path() if world > 1 else None,
                        tokens_per_frame=math.ceil(math.ceil(args.height / 8) / 2) * math.ceil(math.ceil(args.width / 8) / 2),
                        conditions="encoded once per sample (cache), inpaint encoder once per sample"))), flush=True)
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
