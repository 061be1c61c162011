"""TEST INFRASTRUCTURE — CPU restatement (plain PyTorch, fp32) of the reference's STDiT3 denoising
forward, written directly over a state dict in the reference's own order of operations (including
its rearranges, its doubled cross-view batch and its quirks).  It is the oracle the CUDA path is
checked against where the real reference (/root/reference) cannot travel, and the timed CPU
baseline of bench.py.  It is PINNED: tests/test_oracle_cpu.py checks it against the golden vectors
produced by the real reference (tests/golden, oracle/make_golden.py) for every fixture.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this file; the product never does.

Citations: "stdit3.py" = magicdrivedit/models/magicdrive/magicdrive_stdit3.py, "blocks.py" =
magicdrivedit/models/layers/blocks.py, "embedder.py" = magicdrivedit/models/magicdrive/embedder.py.
"""
import math

import torch
import torch.nn.functional as F
from einops import rearrange, repeat


class P:
    """Prefix view over a flat state dict."""

    def __init__(self, sd, prefix=""):
        self.sd, self.prefix = sd, prefix

    def __getitem__(self, k):
        return self.sd[self.prefix + k]

    def has(self, k):
        return (self.prefix + k) in self.sd

    def sub(self, k):
        return P(self.sd, self.prefix + k + ".")

    def get(self, k):
        return self.sd.get(self.prefix + k)


LORA_SCALING = 2.0          # lora_alpha / lora_rank of the BrushNet variant (set by forward from cfg)


def linear(p, name, x):
    r = F.linear(x, p[name + ".weight"], p.get(name + ".bias"))
    if p.has(name + ".lora_A"):                       # lora_forward, stdit3.py:1452-1462
        r = r + ((x @ p[name + ".lora_A"].T) @ p[name + ".lora_B"].T) * LORA_SCALING
    return r


def layernorm(x):                                   # blocks.py:71-80 (eps 1e-6, no affine)
    return F.layer_norm(x, (x.shape[-1],), eps=1e-6)


def t2i_modulate(x, shift, scale):                  # blocks.py:92-93
    return x * (1 + scale) + shift


def rmsnorm(x, w, eps=1e-6):                        # blocks.py:54-68
    dt = x.dtype
    x = x.to(torch.float32)
    x = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps)
    return w * x.to(dt)


def rope(freqs, t):                                 # rotary-embedding-torch 0.5.3 (see oracle/shims)
    pos = torch.arange(t.shape[-2], device=t.device, dtype=t.dtype)
    ang = torch.einsum("..., f -> ... f", pos.type(freqs.dtype), freqs)
    ang = repeat(ang, "... n -> ... (n r)", r=2)
    x1, x2 = rearrange(t, "... (d r) -> ... d r", r=2).unbind(-1)
    rot = rearrange(torch.stack((-x2, x1), -1), "... d r -> ... (d r)")
    return t * ang.cos() + rot * ang.sin()


def naive_attention(q, k, v, scale):                # blocks.py:257-267 / :496-504 / :727-742
    dtype = q.dtype
    # same arithmetic per (batch, head); the batch is walked in slices so the [B,H,Lq,Lk] fp32 score
    # tensor the reference materialises at once (120 GB at S=5600) stays under ~2 GB on the host
    per = q.shape[1] * q.shape[2] * k.shape[2] * 4
    step = max(1, int(2e9 // max(per, 1)))
    outs = []
    for i in range(0, q.shape[0], step):
        qi, ki, vi = q[i:i + step], k[i:i + step], v[i:i + step]
        attn = ((qi * scale) @ ki.transpose(-2, -1)).to(torch.float32).softmax(dim=-1).to(dtype)
        outs.append(attn @ vi)
    return outs[0] if len(outs) == 1 else torch.cat(outs, 0)


def mlp(p, x):                                      # timm Mlp: fc1 -> GELU(tanh) -> fc2
    return linear(p, "fc2", F.gelu(linear(p, "fc1", x), approximate="tanh"))


def attention_temporal(p, x, num_heads, freqs):     # Attention.forward, blocks.py:198-278
    B, N, C = x.shape
    D = C // num_heads
    qkv = linear(p, "qkv", x).view(B, N, 3, num_heads, D).permute(2, 0, 3, 1, 4)
    q, k, v = qkv.unbind(0)
    if p.has("q_norm.weight"):
        q, k = rmsnorm(q, p["q_norm.weight"]), rmsnorm(k, p["k_norm.weight"])
    if freqs is not None:
        q, k = rope(freqs, q), rope(freqs, k)
    x = naive_attention(q, k, v, D ** -0.5).transpose(1, 2).reshape(B, N, C)
    return linear(p, "proj", x)


def attention_mha(p, x, cond, num_heads):           # MultiHeadAttention.forward, blocks.py:425-514
    B, N, C = x.shape
    D = C // num_heads
    cond = x if cond is None else cond
    Nc = cond.shape[1]
    W, b = p["qkv.weight"], p.get("qkv.bias")
    q = F.linear(x, W[:C], None if b is None else b[:C]).view(B, N, num_heads, D)
    kv = F.linear(cond, W[C:], None if b is None else b[C:]).view(B, Nc, 2, num_heads, D)
    k, v = kv.unbind(2)
    q, k = rmsnorm(q, p["q_norm.weight"]), rmsnorm(k, p["k_norm.weight"])
    q, k, v = (t.permute(0, 2, 1, 3) for t in (q, k, v))
    x = naive_attention(q, k, v, D ** -0.5).transpose(1, 2).reshape(B, N, C)
    return linear(p, "proj", x)


def cross_attention(p, x, cond, num_heads):         # MultiHeadCrossAttention.forward, blocks.py:678-747
    B, N, C = x.shape
    D = C // num_heads
    q = linear(p, "q_linear", x).view(B, N, num_heads, D).permute(0, 2, 1, 3)
    kv = linear(p, "kv_linear", cond).view(B, -1, 2, num_heads, D)
    k, v = kv.unbind(2)
    x = naive_attention(q, k.permute(0, 2, 1, 3), v.permute(0, 2, 1, 3), D ** -0.5)
    return linear(p, "proj", x.transpose(1, 2).reshape(B, -1, C))


def t_mask_select(x_mask, x, masked_x, T, S):       # stdit3.py:148-156
    x = rearrange(x, "B (T S) C -> B T S C", T=T, S=S)
    masked_x = rearrange(masked_x, "B (T S) C -> B T S C", T=T, S=S)
    return rearrange(torch.where(x_mask[:, :, None, None], x, masked_x), "B T S C -> B (T S) C")


def block_forward(p, x, y, t, x_mask, t0, T, S, NC, mv_order_map, num_heads, temporal, freqs):
    """MultiViewSTDiT3Block.forward, stdit3.py:192-358."""
    B, N, C = x.shape
    b = B // NC

    def six(tv):
        return repeat(p["scale_shift_table"][None] + tv.reshape(b, 6, -1), "b ... -> (b NC) ...", NC=NC).chunk(6, dim=1)
    sh_a, sc_a, g_a, sh_m, sc_m, g_m = six(t)
    if x_mask is not None:
        sh_a0, sc_a0, g_a0, sh_m0, sc_m0, g_m0 = six(t0)

    def sel(a, a0):
        return t_mask_select(x_mask, a, a0, T, S) if x_mask is not None else a

    x_m = sel(t2i_modulate(layernorm(x), sh_a, sc_a),
              t2i_modulate(layernorm(x), sh_a0, sc_a0) if x_mask is not None else None)
    if temporal:
        x_m = rearrange(x_m, "B (T S) C -> (B S) T C", T=T, S=S)
        x_m = attention_temporal(p.sub("attn"), x_m, num_heads, freqs)
        x_m = rearrange(x_m, "(B S) T C -> B (T S) C", T=T, S=S)
    else:
        x_m = rearrange(x_m, "B (T S) C -> (B T) S C", T=T, S=S)
        x_m = attention_mha(p.sub("attn"), x_m, None, num_heads)
        x_m = rearrange(x_m, "(B T) S C -> B (T S) C", T=T, S=S)
    x = x + sel(g_a * x_m, g_a0 * x_m if x_mask is not None else None)

    if p.has("cross_attn.q_linear.weight"):         # stdit3.py:264-280
        if y.shape[1] == 1:
            x_c = cross_attention(p.sub("cross_attn"), x, y[:, 0], num_heads)
        else:
            x_c = rearrange(x, "B (T S) C -> (B T) S C", T=T, S=S)
            y_c = rearrange(y, "B T L C -> (B T) L C", T=T)
            x_c = cross_attention(p.sub("cross_attn"), x_c, y_c, num_heads)
            x_c = rearrange(x_c, "(B T) S C -> B (T S) C", T=T, S=S)
        x = x + x_c

    if p.has("cross_view_attn.qkv.weight"):         # stdit3.py:285-332
        def three(tv):
            return repeat(p["scale_shift_table_mva"][None] + tv[:, :3].reshape(b, 3, -1),
                          "b ... -> (b NC) ...", NC=NC).chunk(3, dim=1)
        sh_v, sc_v, g_v = three(t)
        if x_mask is not None:
            sh_v0, sc_v0, g_v0 = three(t0)
        x_v = sel(t2i_modulate(layernorm(x), sh_v, sc_v),
                  t2i_modulate(layernorm(x), sh_v0, sc_v0) if x_mask is not None else None)
        x_mv = rearrange(x_v, "(B NC) (T S) C -> (B T) NC S C", NC=NC, T=T)
        h_q, h_kv, order = [], [], []
        for target, values in mv_order_map.items():  # stdit3.py:175-187
            for nb in values:
                h_q.append(x_mv[:, target])
                h_kv.append(x_mv[:, nb])
                order += [target] * len(x_mv)
        raw = attention_mha(p.sub("cross_view_attn"), torch.cat(h_q, 0), torch.cat(h_kv, 0), num_heads)
        order = torch.LongTensor(order)
        out = torch.zeros_like(x_mv)
        for cam in range(NC):                        # stdit3.py:311-320
            out[:, cam] = rearrange(raw[order == cam], "(n b) ... -> b n ...", b=b * T).sum(dim=1)
        out = rearrange(out, "(B T) NC S C -> (B NC) (T S) C", T=T)
        x = x + linear(p, "mva_proj", sel(g_v * out, g_v0 * out if x_mask is not None else None))

    x_m = sel(t2i_modulate(layernorm(x), sh_m, sc_m),
              t2i_modulate(layernorm(x), sh_m0, sc_m0) if x_mask is not None else None)
    x_m = mlp(p.sub("mlp"), x_m)
    x = x + sel(g_m * x_m, g_m0 * x_m if x_mask is not None else None)
    if p.has("after_proj.weight"):
        return x, linear(p, "after_proj", x)
    return x


# ---------------------------------------------------------------------------------------------
# embedders
# ---------------------------------------------------------------------------------------------
def timestep_embedding(t, dim=256, max_period=10000):       # blocks.py:926-946
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(half, dtype=torch.float32) / half).to(t.device)
    args = t[:, None].float() * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


def t_embed(p, t, dtype):                                    # TimestepEmbedder / SizeEmbedder
    h = timestep_embedding(t).to(dtype)
    return linear(p, "mlp.2", F.silu(linear(p, "mlp.0", h)))


def fourier(x, n_freq=4):                                    # embedder.py:487-532
    out = [x]
    for f in 2.0 ** torch.linspace(0.0, n_freq - 1, steps=n_freq):
        out += [torch.sin(x * f), torch.cos(x * f)]
    return torch.cat(out, -1)


def cog_temp_down(x):                                        # embedder.py:117-141
    B, T, N, D = x.shape
    x = x.permute(0, 2, 3, 1).reshape(B * N, D, T)
    if T % 2 == 1:
        first, rest = x[..., 0], x[..., 1:]
        if rest.shape[-1] > 0:
            rest = F.avg_pool1d(rest, kernel_size=2, stride=2)
        x = torch.cat([first[..., None], rest], dim=-1)
    else:
        x = F.avg_pool1d(x, kernel_size=2, stride=2)
    return x.reshape(B, N, D, x.shape[-1]).permute(0, 3, 1, 2)


def downsample(x, tdf):
    if tdf == -1:
        return x.mean(dim=1, keepdim=True)
    if tdf == 4.5:
        return cog_temp_down(cog_temp_down(x))
    return x


def temporal_layer(p, x, num_heads):                         # embedder.py:455-472 / :646-663
    sh1, sc1, g1, sh2, sc2, g2 = p["scale_shift_table"][None].chunk(6, dim=1)
    x = x + g1 * attention_temporal(p.sub("attn"), t2i_modulate(layernorm(x), sh1, sc1), num_heads, p["rope.freqs"])
    return x + g2 * mlp(p.sub("mlp"), t2i_modulate(layernorm(x), sh2, sc2))


def cam_tokens(p, param, mask):                              # CamEmbedder.embed_cam, embedder.py:549-571
    if param.shape[1] == 4:
        param = param[:, :-1]
    bs = param.shape[0]
    param = torch.where((mask > 0)[:, None, None], param, p["uncond_cam"][None])
    emb = fourier(rearrange(param, "b d c -> (b c) d"))
    return linear(p, "emb2token", rearrange(emb, "(b c) d -> b (c d)", b=bs))


def encode_cam(p, cam, drop_mask, NC, temporal, heads, tdf):  # stdit3.py:909-916
    B, T, S = cam.shape[:3]
    mask = repeat(drop_mask, "b T -> (b NC T S)", NC=NC, S=S)
    tok = cam_tokens(p, rearrange(cam, "B T S ... -> (B T S) ..."), mask)
    if not temporal:
        return linear(p, "after_proj", tok)
    x = temporal_layer(p, rearrange(tok, "(b T S) d -> (b S) T d", T=T, S=S), heads)
    x = rearrange(x, "(b S) T d -> b T S d", S=S, T=T)
    return downsample(linear(p, "final_proj", x), tdf)


def encode_box(p, bbox, drop_mask, heads, tdf):              # stdit3.py:882-907 + embedder.py:302-352,432-483
    B, T, N = bbox["classes"].shape
    m = bbox["masks"]
    drop = repeat(drop_mask, "B T -> B T S", S=N)
    null_mask = torch.ones_like(m)
    null_mask[m == 0] = 0
    mask = torch.ones_like(m)
    mask[m == -1] = 0
    mask[torch.logical_and(m == 1, drop == 0)] = 0
    ref = p["null_pos_feature"]
    mask = mask.flatten().unsqueeze(-1).type_as(ref)
    null_mask = null_mask.flatten().unsqueeze(-1).type_as(ref)
    pts = rearrange(bbox["bboxes"], "b t n ... -> (b t n) ...")
    pos = fourier(pts).reshape(pts.shape[0], -1).type_as(ref)
    pos = pos * null_mask + p["null_pos_feature"][None] * (1 - null_mask)
    pos = pos * mask + p["mask_pos_feature"][None] * (1 - mask)
    idx = bbox["classes"].flatten().long()
    cls = p["_class_tokens"][idx]
    mu, logvar = torch.split(p["mean_var"][idx], 1, dim=1)
    lat = rearrange(bbox["box_latent"], "b t n ... -> (b t n) ...")
    cls = cls + (lat * torch.exp(0.5 * logvar) + mu)
    cls = cls * null_mask + p["null_class_feature"][None] * (1 - null_mask)
    cls = cls * mask + p["mask_class_feature"][None] * (1 - mask)
    emb = torch.cat([F.silu(linear(p, "bbox_proj", pos)), cls], dim=-1)
    emb = linear(p, "second_linear.4", F.silu(linear(p, "second_linear.2", F.silu(linear(p, "second_linear.0", emb)))))
    x = temporal_layer(p, rearrange(emb, "(b t n) d -> (b n) t d", t=T, n=N), heads)
    x = rearrange(x, "(b n) t d -> b t n d", b=B, n=N)
    return downsample(linear(p, "final_proj", x), tdf)


def conv(p, name, x, **kw):
    return F.conv2d(x, p[name + ".weight"], p[name + ".bias"], **kw)


def encode_map(p_all, maps, NC, h_pad, x_shape):              # stdit3.py:963-1007
    p = p_all.sub("controlnet_cond_embedder")
    b, T = maps.shape[:2]
    h = F.silu(conv(p, "conv_in", rearrange(maps, "b T ... -> (b T) ..."), padding=1))   # embedder.py:709-794
    n_blk = len([k for k in p_all.sd if k.startswith(p.prefix + "blocks.") and k.endswith(".weight")])
    for i in range(n_blk):
        last2 = i >= n_blk - 2
        if i % 2 == 0:
            h = conv(p, f"blocks.{i}", h, padding=(2, 1) if last2 else 1)
        else:
            h = conv(p, f"blocks.{i}", h, padding=(2, 1), stride=(2, 1) if last2 else 2)
        h = F.silu(h)
    h = conv(p, "conv_out", h, padding=1)
    h = rearrange(h, "(b T) C ... -> b C T ...", T=T)
    pt = p_all.sub("controlnet_cond_embedder_temp")
    for i in (1, 3):                                          # embedder.py:41-95, :816-821
        h = F.pad(h, (1, 0, 1, 0))
        bs, ch, fr, hh, ww = h.shape
        x = cog_temp_down(h.permute(0, 2, 3, 4, 1).reshape(bs, fr, hh * ww, ch))
        x = x.reshape(bs, x.shape[1], hh, ww, ch).permute(0, 4, 1, 2, 3)
        x = F.pad(x, (0, 1, 0, 1))
        bs, ch, fr, hh, ww = x.shape
        x = conv(pt, f"conv_blocks.{i}.conv", x.permute(0, 2, 1, 3, 4).reshape(bs * fr, ch, hh, ww))
        h = x.reshape(bs, fr, *x.shape[1:]).permute(0, 2, 1, 3, 4)
    if h.shape[-3:] != x_shape[-3:]:
        h = F.interpolate(h, x_shape[-3:])
    if h_pad > 0:
        h = F.pad(h, (0, 0, 0, h_pad * 2))
    pp = p_all.sub("controlnet_cond_patchifier")
    h = F.conv3d(h, pp["proj.weight"], pp["proj.bias"], stride=(1, 2, 2)).flatten(2).transpose(1, 2)
    return repeat(h, "b ... -> (b NC) ...", NC=NC)


def pos_embed(C, h, w, scale, base_size, dtype):             # blocks.py:1067-1116
    half = C // 2
    inv_freq = 1.0 / (10000 ** (torch.arange(0, half, 2).float() / half))
    gh, gw = torch.arange(h) / scale, torch.arange(w) / scale
    gh, gw = gh * (base_size / h), gw * (base_size / w)
    gh, gw = torch.meshgrid(gw, gh, indexing="ij")
    gh, gw = gh.t().reshape(-1), gw.t().reshape(-1)

    def sc(t):
        o = torch.einsum("i,d->id", t, inv_freq)
        return torch.cat((torch.sin(o), torch.cos(o)), dim=-1)
    return torch.concat([sc(gh), sc(gw)], dim=-1).unsqueeze(0).to(dtype)


def patch_embed(p, x):                                       # blocks.py:132-155
    _, _, D, H, W = x.shape
    if W % 2:
        x = F.pad(x, (0, 1))
    if H % 2:
        x = F.pad(x, (0, 0, 0, 1))
    return F.conv3d(x, p["proj.weight"], p["proj.bias"], stride=(1, 2, 2)).flatten(2).transpose(1, 2)


def shallow_encoder(p, x, chunk_size=64):                    # ShallowEncoder.forward, stdit3.py:1284-1311
    B_NC, _, T, _, _ = x.shape
    x = rearrange(x, "b c t h w -> (b t) c h w")
    chunks = []
    for i in range(0, x.shape[0], chunk_size):
        ch = x[i:i + chunk_size]
        ch = F.relu(conv(p, "conv1", ch, stride=2, padding=1))
        ch = F.relu(conv(p, "conv2", ch, stride=2, padding=1))
        ch = F.relu(conv(p, "conv3", ch, stride=2, padding=1))
        chunks.append(conv(p, "conv4", ch, stride=1, padding=1))
    x = rearrange(torch.cat(chunks, dim=0), "(b t) c h w -> b c t h w", b=B_NC, t=T)
    return F.conv3d(x, p["temporal_conv.weight"], p["temporal_conv.bias"], stride=(4, 1, 1), padding=(2, 0, 0))


# ---------------------------------------------------------------------------------------------
# the denoiser forward, stdit3.py:1021-1218 (BrushNet variant with x_inpaint / mask_inpaint: :1466-1867)
# ---------------------------------------------------------------------------------------------
@torch.no_grad()
def forward(sd, cfg, x, timestep, y, maps, bbox, cams, rel_pos, fps, height, width, mv_order_map=None,
            mask=None, x_mask=None, max_layers=None, timers=None, x_inpaint=None, mask_inpaint=None,
            timestep_inpaint=None, num_timesteps=None, noise_inpaint_encoded=None, **_):
    """sd: state dict (fp32 tensors); cfg: constructor kwargs (synth.model_kwargs).
    ``max_layers`` truncates the block stack; ``timers`` collects wall time per block kind (bench.py)."""
    global LORA_SCALING
    LORA_SCALING = cfg.get("lora_alpha", 32) / cfg.get("lora_rank", 16)
    p = P(sd)
    C, NH = cfg["hidden_size"], cfg["num_heads"]
    depth, cdepth = cfg["depth"], cfg["control_depth"]
    if max_layers is not None:
        depth, cdepth = min(depth, max_layers), min(cdepth, max_layers)
    emb_heads = cfg["bbox_embedder_param"]["num_heads"]
    tdf = cfg["bbox_embedder_param"]["time_downsample_factor"]
    dtype = sd["x_embedder.proj.weight"].dtype
    B, real_T = x.size(0), rel_pos.size(1)
    drop_cond = torch.ones((B,), dtype=x.dtype)
    drop_frame = torch.ones((B, real_T), dtype=x.dtype)
    NC = len(mv_order_map)
    x = rearrange(x.to(dtype), "B (C NC) T ... -> (B NC) C T ...", NC=NC)
    timestep, y = timestep.to(dtype), y.to(dtype)
    enc = mi = None
    if x_inpaint is not None:                                # stdit3.py:1504-1524
        xi = rearrange(x_inpaint.to(dtype), "B (C NC) T ... -> (B NC) C T ...", NC=NC)
        mi = rearrange(mask_inpaint.to(dtype), "B (C NC) T ... -> (B NC) C T ...", NC=NC)
        enc = shallow_encoder(p.sub("shallow_encoder"), xi)
        if timestep_inpaint is not None:                     # SDE variant: add_noise, stdit3.py:1976-1994, :2099-2111
            nz = rearrange(noise_inpaint_encoded.to(dtype), "B (C NC) T ... -> (B NC) C T ...", NC=NC)
            assert nz.shape == enc.shape
            tp = 1 - repeat(timestep_inpaint, "b -> (b NC)", NC=NC).float() / num_timesteps
            tp = tp[:, None, None, None, None].repeat(1, *nz.shape[1:])
            enc = (tp * enc + (1 - tp) * nz).to(dtype)
        mi = F.interpolate(mi, size=x.shape[-3:], mode="trilinear", align_corners=False)
        assert x.shape[-3:] == enc.shape[-3:] == mi.shape[-3:]
    _, _, Tx, Hx, Wx = x.size()
    x_in_shape = x.shape
    T, H, W = Tx, -(-Hx // 2), -(-Wx // 2)
    S = H * W
    h_pad = 0
    q = cfg.get("force_pad_h_for_sp_size")
    if q is not None and S % q != 0:                         # stdit3.py:1061-1063
        h_pad = q - H % q
    if h_pad > 0:
        x = F.pad(x, (0, 0, 0, h_pad * 2))
        if enc is not None:                                  # stdit3.py:1586-1587
            enc, mi = F.pad(enc, (0, 0, 0, h_pad * 2)), F.pad(mi, (0, 0, 0, h_pad * 2))
        H += h_pad
        S = H * W
    base_size = round(S ** 0.5)
    scale = (height[0].item() * width[0].item()) ** 0.5 / cfg.get("input_sq_size", 512)
    pos = pos_embed(C, H, W, scale, base_size, dtype)

    fps_ = fps.unsqueeze(1)
    if fps_.shape[0] != B:                                   # SizeEmbedder.forward, blocks.py:999-1003
        fps_ = fps_.repeat(B // fps_.shape[0], 1)
    fps_e = t_embed(p.sub("fps_embedder"), fps_.reshape(-1), dtype).reshape(B, -1)
    t = t_embed(p.sub("t_embedder"), timestep, dtype) + fps_e
    t_mlp = linear(p, "t_block.1", F.silu(t))
    t0 = t0_mlp = None
    if x_mask is not None:
        t0 = t_embed(p.sub("t_embedder"), torch.zeros_like(timestep), dtype) + fps_e
        t0_mlp = linear(p, "t_block.1", F.silu(t0))

    tb_mlp, tb0_mlp = t_mlp, t0_mlp                          # time vectors of the brushnet blocks
    if timestep_inpaint is not None:                         # SDE variant, stdit3.py:2215-2232
        t_in = t_embed(p.sub("t_embedder"), timestep_inpaint, dtype) + fps_e
        tb_mlp = linear(p, "t_combine_block.1",
                        F.silu(torch.cat([t_mlp, linear(p, "t_inpaint_block.1", F.silu(t_in))], dim=-1)))
        if x_mask is not None:
            tb0_mlp = linear(p, "t_combine_block.1",
                             F.silu(torch.cat([t0_mlp, linear(p, "t_inpaint_block.1", F.silu(t0))], dim=-1)))

    # condition sequence, stdit3.py:918-961
    yt = mlp(p.sub("y_embedder.y_proj"), y)                   # [B,1,L,C]
    if mask is not None:
        m = mask.repeat(yt.shape[0] // mask.shape[0], 1) if mask.shape[0] != yt.shape[0] else mask
        max_len = int(min(max(i + 1 for i in m.sum(dim=1).tolist()), yt.shape[2]))
        yt = yt.squeeze(1)[:, :max_len]
    else:
        yt = yt.squeeze(1)
    yt = repeat(yt, "b ... -> (b NC) ...", NC=NC)
    base = sd["base_token"][None, None, None]
    drop_box = repeat(torch.logical_and(drop_cond[:, None], drop_frame), "b ... -> (b NC) ...", NC=NC)
    box_emb = base + encode_box(p.sub("bbox_embedder"), bbox, drop_box, emb_heads, tdf)
    cam_emb = encode_cam(p.sub("camera_embedder"), cams[:, 0:1], repeat(drop_cond, "b -> b T", T=1), NC, False, 0, 0)
    frame_emb = base + encode_cam(p.sub("frame_embedder"), rel_pos, drop_frame, NC, True, emb_heads, tdf)
    cam_emb = base + rearrange(cam_emb, "(B 1 S) ... -> B 1 S ...", S=cams.shape[2])
    cam_emb = repeat(cam_emb, "B 1 S ... -> B T S ...", T=frame_emb.shape[1])
    yt = repeat(yt, "B ... -> B T ...", T=frame_emb.shape[1])
    cond = torch.cat([frame_emb, cam_emb, yt, box_emb], dim=2)
    if cond.shape[1] != T and cond.shape[1] > 1:
        L_ = cond.shape[2]
        cond = rearrange(F.interpolate(rearrange(cond, "B T L D -> B (L D) T"), T), "B (L D) T -> B T L D", L=L_)

    c = encode_map(p, maps, NC, h_pad, x_in_shape)
    c = rearrange(c, "B (T S) C -> B T S C", T=T)
    x_b = rearrange(patch_embed(p.sub("x_embedder"), x), "B (T S) C -> B T S C", T=T, S=S) + pos
    xe = "x_control_embedder" if p.has("x_control_embedder.proj.weight") else "x_embedder"
    x_c = rearrange(patch_embed(p.sub(xe), x), "B (T S) C -> B T S C", T=T, S=S) + pos
    c = x_c + linear(p, "before_proj", c)
    x_inp = None
    if enc is not None:                                      # stdit3.py:1648-1654
        x_inp = patch_embed(p.sub("x_brushnet_embedder"), torch.cat([x, enc, mi], dim=1))
        x_inp = rearrange(rearrange(x_inp, "B (T S) C -> B T S C", T=T, S=S) + pos, "B T S C -> B (T S) C")
    x = rearrange(x_b, "B T S C -> B (T S) C")
    c = rearrange(c, "B T S C -> B (T S) C")
    if x_mask is not None:
        x_mask = repeat(x_mask, "b ... -> (b NC) ...", NC=NC)
    freqs = sd["rope.freqs"]
    has_t = p.has("base_blocks_t.0.attn.qkv.weight")
    has_ct = p.has("control_blocks_t.0.attn.qkv.weight")

    def blk(name, i, h, temporal):
        import time
        t_ = time.perf_counter()
        brush = name.startswith("brushnet")
        r = block_forward(p.sub(f"{name}.{i}"), h, cond, tb_mlp if brush else t_mlp, x_mask,
                          tb0_mlp if brush else t0_mlp, T, S, NC, mv_order_map, NH, temporal,
                          freqs if temporal else None)
        if timers is not None:                               # bench.py: wall seconds per block kind
            timers[name] = timers.get(name, 0.0) + time.perf_counter() - t_
        return r
    has_bt = p.has("brushnet_blocks_t.0.attn.qkv.weight")
    for i in range(depth):                                   # stdit3.py:1173-1198 / BrushNet :1684-1850
        x = blk("base_blocks_s", i, x, False)
        skip = 0
        if i < cdepth:
            c, skip = blk("control_blocks_s", i, c, False)
        if x_inp is not None:
            x_inp, skip_i = blk("brushnet_blocks_s", i, x_inp, False)
            x = x + skip + skip_i if i < cdepth else x + skip_i
        elif i < cdepth:
            x = x + skip
        if has_t:
            x = blk("base_blocks_t", i, x, True)
        if i < cdepth and has_ct:
            c, skip = blk("control_blocks_t", i, c, True)
            x = x + skip
        if x_inp is not None and has_bt and has_t:
            x_inp, skip_i = blk("brushnet_blocks_t", i, x_inp, True)
            x = x + skip_i

    # final layer incl. its t0 quirk (blocks.py:893-905): x_zero is built from the modulated x
    fl = p.sub("final_layer")
    tt = repeat(t, "b d -> (b NC) d", NC=NC)
    shift, scale_ = (fl["scale_shift_table"][None] + tt[:, None]).chunk(2, dim=1)
    x = t2i_modulate(layernorm(x), shift, scale_)
    if x_mask is not None:
        tt0 = repeat(t0, "b d -> (b NC) d", NC=NC)
        shift0, scale0 = (fl["scale_shift_table"][None] + tt0[:, None]).chunk(2, dim=1)
        x = t_mask_select(x_mask, x, t2i_modulate(layernorm(x), shift0, scale0), T, S)
    x = linear(fl, "linear", x)
    Co = x.shape[-1] // 4
    x = rearrange(x, "B (N_t N_h N_w) (T_p H_p W_p C_out) -> B C_out (N_t T_p) (N_h H_p) (N_w W_p)",
                  N_t=T, N_h=H, N_w=W, T_p=1, H_p=2, W_p=2, C_out=Co)[:, :, :Tx, :Hx, :Wx]
    return rearrange(x.to(torch.float32), "(B NC) C T ... -> B (C NC) T ...", NC=NC)
