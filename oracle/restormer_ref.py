"""ORACLE (test infrastructure, not product code): CPU fp32 restatement of the reference Restormer forward.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
It restates AgenticIR/executor/defocus_deblurring/tools/Restormer/basicsr/models/archs/restormer_arch.py as
pure functions over a state_dict (the reference builds nn.Modules). Pinned against the reference module itself,
imported unmodified in the build container, by tools/make_golden.py -> tests/golden/restormer_*.npz.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def layer_norm(x, weight, bias):
    """restormer_arch.py:25-70. x [B,C,H,W]; normalises over C per pixel, population variance, eps 1e-5.
    bias=None is BiasFree_LayerNorm (:37-39): no mean subtraction in the numerator."""
    t = x.permute(0, 2, 3, 1)  # to_3d (:18-19) keeps channel last
    sigma = t.var(-1, keepdim=True, unbiased=False)
    if bias is None:
        y = t / torch.sqrt(sigma + 1e-5) * weight
    else:
        mu = t.mean(-1, keepdim=True)
        y = (t - mu) / torch.sqrt(sigma + 1e-5) * weight + bias
    return y.permute(0, 3, 1, 2)


def mdta(x, sd, p, heads):
    """Attention.forward, restormer_arch.py:111-132."""
    b, c, h, w = x.shape
    qkv = F.conv2d(x, sd[f"{p}.qkv.weight"])
    qkv = F.conv2d(qkv, sd[f"{p}.qkv_dwconv.weight"], padding=1, groups=3 * c)
    q, k, v = qkv.chunk(3, dim=1)
    q = q.reshape(b, heads, c // heads, h * w)
    k = k.reshape(b, heads, c // heads, h * w)
    v = v.reshape(b, heads, c // heads, h * w)
    q = F.normalize(q, dim=-1)
    k = F.normalize(k, dim=-1)
    attn = (q @ k.transpose(-2, -1)) * sd[f"{p}.temperature"]
    attn = attn.softmax(dim=-1)
    out = (attn @ v).reshape(b, c, h, w)
    return F.conv2d(out, sd[f"{p}.project_out.weight"])


def gdfn(x, sd, p):
    """FeedForward.forward, restormer_arch.py:88-93."""
    w_in = sd[f"{p}.project_in.weight"]
    x = F.conv2d(x, w_in)
    x = F.conv2d(x, sd[f"{p}.dwconv.weight"], padding=1, groups=w_in.shape[0])
    x1, x2 = x.chunk(2, dim=1)
    return F.conv2d(F.gelu(x1) * x2, sd[f"{p}.project_out.weight"])


def block(x, sd, p, heads):
    """TransformerBlock.forward, restormer_arch.py:146-150."""
    x = x + mdta(layer_norm(x, sd[f"{p}.norm1.body.weight"], sd.get(f"{p}.norm1.body.bias")), sd, f"{p}.attn", heads)
    x = x + gdfn(layer_norm(x, sd[f"{p}.norm2.body.weight"], sd.get(f"{p}.norm2.body.bias")), sd, f"{p}.ffn")
    return x


def _stage(x, sd, name, n, heads):
    for i in range(n):
        x = block(x, sd, f"{name}.{i}", heads)
    return x


def forward(sd, cfg, inp_img):
    """Restormer.forward, restormer_arch.py:245-284 (dual_pixel_task=False branch)."""
    nb, hd = cfg["num_blocks"], cfg["heads"]
    down = lambda t, n: F.pixel_unshuffle(F.conv2d(t, sd[f"{n}.body.0.weight"], padding=1), 2)  # :171-179
    up = lambda t, n: F.pixel_shuffle(F.conv2d(t, sd[f"{n}.body.0.weight"], padding=1), 2)      # :181-189
    e1 = _stage(F.conv2d(inp_img, sd["patch_embed.proj.weight"], padding=1), sd, "encoder_level1", nb[0], hd[0])
    e2 = _stage(down(e1, "down1_2"), sd, "encoder_level2", nb[1], hd[1])
    e3 = _stage(down(e2, "down2_3"), sd, "encoder_level3", nb[2], hd[2])
    lat = _stage(down(e3, "down3_4"), sd, "latent", nb[3], hd[3])
    d3 = torch.cat([up(lat, "up4_3"), e3], 1)
    d3 = _stage(F.conv2d(d3, sd["reduce_chan_level3.weight"]), sd, "decoder_level3", nb[2], hd[2])
    d2 = torch.cat([up(d3, "up3_2"), e2], 1)
    d2 = _stage(F.conv2d(d2, sd["reduce_chan_level2.weight"]), sd, "decoder_level2", nb[1], hd[1])
    d1 = torch.cat([up(d2, "up2_1"), e1], 1)
    d1 = _stage(d1, sd, "decoder_level1", nb[0], hd[0])
    d1 = _stage(d1, sd, "refinement", cfg["num_refinement_blocks"], hd[0])
    return F.conv2d(d1, sd["output.weight"], padding=1) + inp_img


class Model:
    """Callable with the signature of the reference nn.Module: model(x[B,c,H,W]) -> [B,c,H,W]."""

    def __init__(self, sd, cfg):
        self.sd = {k: v.float() for k, v in sd.items()}
        self.cfg = cfg

    @torch.no_grad()
    def __call__(self, x):
        return forward(self.sd, self.cfg, x)
