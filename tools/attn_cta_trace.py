"""Life line of EVERY CTA of one attention launch (TRACE build of the fast kernel, global timer): how long a CTA lives,
how much of that is before its first score tile and after its last probability tile, and how long an SM sits between
two CTAs.  Fits  t_cta = fixed + per_tile * n_tiles  over several sequence lengths."""
import ctypes, os, sys
os.environ["MDV2_ATTN_PERSIST"] = "0"   # the life line is a property of the one-CTA-per-item kernel
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from magicdrive_v2_b200 import ops

STEPS, EV, CTA_EV = 96, 12, 8
lib = ops._lib
lib.mdv2_debug_attn_trace.argtypes = [ctypes.c_void_p]
lib.mdv2_debug_attn_trace.restype = None
H, D = 16, 72
C = H * D


def run(n_seq, L, n_sets=1):
    qkv = torch.randn(n_seq * L, 3 * C, device="cuda", dtype=torch.bfloat16)
    out = torch.empty(n_seq * L, C, device="cuda", dtype=torch.bfloat16)
    kv_map = None
    if n_sets == 2:
        kv_map = torch.stack([(torch.arange(n_seq) + 1) % n_seq, (torch.arange(n_seq) + n_seq - 1) % n_seq], 1).int().contiguous().cuda()
    n_cta = ((L + 255) // 256) * H * n_seq
    buf = torch.zeros(20 * STEPS * EV + n_cta * CTA_EV, dtype=torch.int64, device="cuda")
    call = lambda: ops.attention(qkv[:, :C], qkv[:, C:2 * C], qkv[:, 2 * C:], out, n_seq, L, L, H, D, D ** -0.5, kv_map=kv_map)
    call(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); call(); e1.record(); torch.cuda.synchronize()
    ms_plain = e0.elapsed_time(e1)
    lib.mdv2_debug_attn_trace(buf.data_ptr())
    call(); torch.cuda.synchronize()
    lib.mdv2_debug_attn_trace(None)
    t = buf[20 * STEPS * EV:].view(n_cta, CTA_EV).cpu().double()
    life = (t[:, 5] - t[:, 0]) / 1e3
    head = (t[:, 2] - t[:, 0]) / 1e3
    tail = (t[:, 5] - t[:, 3]) / 1e3
    otail = (t[:, 5] - t[:, 4]) / 1e3
    main = (t[:, 3] - t[:, 2]) / 1e3
    gaps = []
    sm = t[:, 1].long()
    for s in sm.unique().tolist():
        rows = t[sm == s]
        rows = rows[rows[:, 0].argsort()]
        if len(rows) > 1:
            gaps.append((rows[1:, 0] - rows[:-1, 5]) / 1e3)
    gaps = torch.cat(gaps)
    span = (t[:, 5].max() - t[:, 0].min()) / 1e6
    full = (torch.arange(n_cta) % ((L + 255) // 256)) != ((L + 255) // 256 - 1)    # CTAs with two whole query tiles
    print(f"n_seq={n_seq:3d} L={L:5d} sets={n_sets} tiles/pass={(L + 127) // 128:3d}: plain {ms_plain:7.3f} ms, traced span {span:7.3f} ms, "
          f"{n_cta} CTAs on {len(sm.unique())} SMs | per CTA (us, median): life {life[full].median():7.2f}  "
          f"start->first S {head[full].median():5.2f}  first S->last P {main[full].median():7.2f}  last P->end {tail[full].median():5.2f} "
          f"(O ready->end {otail[full].median():5.2f})  gap between CTAs on an SM {gaps.median():5.2f} (mean {gaps.mean():5.2f})", flush=True)


if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    for n_seq, L in ((480, 512), (240, 1024), (204, 1600), (120, 2048), (60, 4096), (60, 5600), (30, 8192)):
        run(n_seq, L)
    run(60, 5600, 2)
    run(204, 1600, 2)
