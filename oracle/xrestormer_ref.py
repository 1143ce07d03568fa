"""ORACLE (test infrastructure, not product code): CPU fp32 restatement of the reference X-Restormer forward
(SURVEY.md §8f row 3).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
It restates AgenticIR/executor/dehazing/tools/X-Restormer/xrestormer/archs/xrestormer_arch.py as pure functions over
a state_dict. Pinned against the reference module itself (imported in the build container with shims for its
basicsr registry / torchstat imports) by tools/make_golden_full.py -> tests/golden/xrestormer_*.npz.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .restormer_ref import gdfn, layer_norm, mdta


def rel_pos_bias(q, rel_height, rel_width, block: int, rel_size: int):
    """RelPosEmb.forward (xrestormer_arch.py:91-103 with rel_to_abs / relative_logits_1d :31-62) without the
    pad-and-reshape skewing: for the query at (x, y) of the block and the key at (i, j) of the rel_size x rel_size
    window the logit is q . rel_width[j - y + rel_size - 1] + q . rel_height[i - x + rel_size - 1].
    q: [B, block*block, d] (already scaled). Returns [B, block*block, rel_size*rel_size]."""
    B, n, d = q.shape
    r = rel_size
    lw = q @ rel_width.t()   # [B, n, 2r-1]
    lh = q @ rel_height.t()
    xs = torch.arange(block).repeat_interleave(block)       # query row x of token (x y)
    ys = torch.arange(block).repeat(block)                  # query column y
    ki = torch.arange(r).repeat_interleave(r)               # key row i of (i j)
    kj = torch.arange(r).repeat(r)                          # key column j
    idx_w = (kj[None, :] - ys[:, None] + r - 1)             # [n, r*r]
    idx_h = (ki[None, :] - xs[:, None] + r - 1)
    return torch.gather(lw, 2, idx_w.expand(B, -1, -1)) + torch.gather(lh, 2, idx_h.expand(B, -1, -1))


def ocab(x, sd, p, heads, window: int, overlap_ratio: float, dim_head: int):
    """OCAB.forward, xrestormer_arch.py:190-237: queries from ws x ws blocks, keys / values from the
    (ws + int(ws * ratio))-wide windows around them (nn.Unfold, stride ws, ZERO padding), relative-position logits
    from the scaled queries, softmax over all window positions including the padded ones."""
    b, c, h, w = x.shape
    ows = int(window * overlap_ratio) + window
    inner = dim_head * heads
    qkv = F.conv2d(x, sd[f"{p}.qkv.weight"])
    qs, ks, vs = qkv.chunk(3, dim=1)
    nh, nw = h // window, w // window
    # 'b c (h p1) (w p2) -> (b h w) (p1 p2) c'
    qs = qs.reshape(b, inner, nh, window, nw, window).permute(0, 2, 4, 3, 5, 1).reshape(b * nh * nw, window * window, inner)
    pad = (ows - window) // 2

    def unfold(t):  # 'b (c j) i -> (b i) j c'
        u = F.unfold(t, kernel_size=(ows, ows), stride=window, padding=pad)        # [b, inner*ows*ows, nh*nw]
        return u.reshape(b, inner, ows * ows, nh * nw).permute(0, 3, 2, 1).reshape(b * nh * nw, ows * ows, inner)

    ks, vs = unfold(ks), unfold(vs)

    def split(t):  # 'b n (head c) -> (b head) n c'
        B, n, _ = t.shape
        return t.reshape(B, n, heads, dim_head).permute(0, 2, 1, 3).reshape(B * heads, n, dim_head)

    qs, ks, vs = split(qs), split(ks), split(vs)
    qs = qs * dim_head ** -0.5
    attn = qs @ ks.transpose(-2, -1)
    attn = attn + rel_pos_bias(qs, sd[f"{p}.rel_pos_emb.rel_height"], sd[f"{p}.rel_pos_emb.rel_width"], window, ows)
    attn = attn.softmax(dim=-1)
    out = attn @ vs                                                                 # [(b h w head), ws*ws, d]
    # '(b h w head) (p1 p2) c -> b (head c) (h p1) (w p2)'
    out = out.reshape(b, nh, nw, heads, window, window, dim_head).permute(0, 3, 6, 1, 4, 2, 5).reshape(b, inner, h, w)
    return F.conv2d(out, sd[f"{p}.project_out.weight"])


def block(x, sd, p, cfg, level):
    """TransformerBlock.forward, xrestormer_arch.py:256-261."""
    ln = lambda t, n: layer_norm(t, sd[f"{p}.{n}.body.weight"], sd.get(f"{p}.{n}.body.bias"))
    x = x + mdta(ln(x, "norm1"), sd, f"{p}.channel_attn", cfg["channel_heads"][level])
    x = x + gdfn(ln(x, "norm2"), sd, f"{p}.channel_ffn")
    x = x + ocab(ln(x, "norm3"), sd, f"{p}.spatial_attn", cfg["spatial_heads"][level], cfg["window_size"],
                 cfg["overlap_ratio"][level], cfg["spatial_dim_head"])
    x = x + gdfn(ln(x, "norm4"), sd, f"{p}.spatial_ffn")
    return x


def _stage(x, sd, name, n, cfg, level):
    for i in range(n):
        x = block(x, sd, f"{name}.{i}", cfg, level)
    return x


def forward(sd, cfg, inp_img):
    """XRestormer.forward, xrestormer_arch.py:382-412."""
    if cfg.get("scale", 1) > 1:  # :383-384
        inp_img = F.interpolate(inp_img, scale_factor=cfg["scale"], mode="bilinear", align_corners=False)
    nb = cfg["num_blocks"]
    down = lambda t, n: F.pixel_unshuffle(F.conv2d(t, sd[f"{n}.body.0.weight"], padding=1), 2)
    up = lambda t, n: F.pixel_shuffle(F.conv2d(t, sd[f"{n}.body.0.weight"], padding=1), 2)
    e1 = _stage(F.conv2d(inp_img, sd["patch_embed.proj.weight"], padding=1), sd, "encoder_level1", nb[0], cfg, 0)
    e2 = _stage(down(e1, "down1_2"), sd, "encoder_level2", nb[1], cfg, 1)
    e3 = _stage(down(e2, "down2_3"), sd, "encoder_level3", nb[2], cfg, 2)
    lat = _stage(down(e3, "down3_4"), sd, "latent", nb[3], cfg, 3)
    d3 = torch.cat([up(lat, "up4_3"), e3], 1)
    d3 = _stage(F.conv2d(d3, sd["reduce_chan_level3.weight"]), sd, "decoder_level3", nb[2], cfg, 2)
    d2 = torch.cat([up(d3, "up3_2"), e2], 1)
    d2 = _stage(F.conv2d(d2, sd["reduce_chan_level2.weight"]), sd, "decoder_level2", nb[1], cfg, 1)
    d1 = torch.cat([up(d2, "up2_1"), e1], 1)
    d1 = _stage(d1, sd, "decoder_level1", nb[0], cfg, 0)
    d1 = _stage(d1, sd, "refinement", cfg["num_refinement_blocks"], cfg, 0)
    return F.conv2d(d1, sd["output.weight"], padding=1) + inp_img


class Model:
    """Callable with the signature of the reference nn.Module: model(x[B,c,H,W]) -> [B,c,H,W]; H, W multiples of 64
    (8 x window 8 at the deepest level; the tool pads to 64, xrestormer_model.py:56-66)."""

    def __init__(self, sd, cfg):
        self.sd = {k: v.float() for k, v in sd.items()}
        self.cfg = cfg

    @torch.no_grad()
    def __call__(self, x):
        return forward(self.sd, self.cfg, x)
