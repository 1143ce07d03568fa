"""ORACLE (test infrastructure, not product code): CPU fp32 restatement of the reference HAT forward
(SURVEY.md §8f row 4).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
Restates AgenticIR/executor/super_resolution/tools/HAT/hat/archs/hat_arch.py as pure functions over a state_dict.
The relative-position indices and the shifted-window mask are closed forms of the registered buffers
(calculate_rpi_sa :882-894, calculate_rpi_oca :896-919, calculate_mask :921-941). Pinned against the reference module
itself (imported in the build container with stand-ins for its basicsr registry / arch_util imports) by
tools/make_golden_full.py -> tests/golden/hat_*.npz.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from .swinir_ref import relative_position_index, shift_mask


def rpi_oca(ws: int, ows: int) -> torch.Tensor:
    """calculate_rpi_oca, hat_arch.py:896-919: for query (qy, qx) of the ws x ws window and key (ky, kx) of the
    ows x ows window: ((ky - qy) + ws - ows + 1) * (ws + ows - 1) + ((kx - qx) + ws - ows + 1).
    The reference's offset (ws - ows + 1, not ws - 1) makes part of these indices NEGATIVE; they index the table from
    its end (PyTorch advanced indexing), which the oracle reproduces by keeping them negative."""
    q = torch.arange(ws * ws)
    k = torch.arange(ows * ows)
    qy, qx = q // ws, q % ws
    ky, kx = k // ows, k % ows
    off = ws - ows + 1
    return (ky[None, :] - qy[:, None] + off) * (ws + ows - 1) + (kx[None, :] - qx[:, None] + off)


def _mlp(x, sd, p):
    """Mlp.forward :88-94."""
    return F.linear(F.gelu(F.linear(x, sd[f"{p}.fc1.weight"], sd[f"{p}.fc1.bias"])), sd[f"{p}.fc2.weight"], sd[f"{p}.fc2.bias"])


def _partition(x, ws):
    """window_partition :97-109. x [b, h, w, c] -> [nw*b, ws*ws, c]."""
    b, h, w, c = x.shape
    return x.view(b, h // ws, ws, w // ws, ws, c).permute(0, 1, 3, 2, 4, 5).reshape(-1, ws * ws, c)


def _reverse(xw, ws, b, h, w):
    """window_reverse :112-126."""
    return xw.view(b, h // ws, w // ws, ws, ws, -1).permute(0, 1, 3, 2, 4, 5).reshape(b, h, w, -1)


def cab(x, sd, p):
    """CAB.forward :73-74 with ChannelAttention.forward :56-58. x [b, c, h, w]."""
    y = F.gelu(F.conv2d(x, sd[f"{p}.cab.0.weight"], sd[f"{p}.cab.0.bias"], padding=1))
    y = F.conv2d(y, sd[f"{p}.cab.2.weight"], sd[f"{p}.cab.2.bias"], padding=1)
    a = y.mean(dim=(2, 3), keepdim=True)                                          # AdaptiveAvgPool2d(1)
    a = F.relu(F.conv2d(a, sd[f"{p}.cab.3.attention.1.weight"], sd[f"{p}.cab.3.attention.1.bias"]))
    a = torch.sigmoid(F.conv2d(a, sd[f"{p}.cab.3.attention.3.weight"], sd[f"{p}.cab.3.attention.3.bias"]))
    return y * a


def window_attention(xw, sd, p, heads, ws, mask):
    """WindowAttention.forward :165-197 (the same arithmetic as SwinIR's, the index passed in as rpi_sa)."""
    b_, n, c = xw.shape
    qkv = F.linear(xw, sd[f"{p}.qkv.weight"], sd[f"{p}.qkv.bias"]).reshape(b_, n, 3, heads, c // heads)
    q, k, v = qkv.permute(2, 0, 3, 1, 4)
    attn = (q * (c // heads) ** -0.5) @ k.transpose(-2, -1)
    bias = sd[f"{p}.relative_position_bias_table"][relative_position_index(ws).reshape(-1)].reshape(n, n, heads)
    attn = attn + bias.permute(2, 0, 1).unsqueeze(0)
    if mask is not None:
        nw = mask.shape[0]
        attn = (attn.view(b_ // nw, nw, heads, n, n) + mask[None, :, None]).view(-1, heads, n, n)
    out = (attn.softmax(-1) @ v).transpose(1, 2).reshape(b_, n, c)
    return F.linear(out, sd[f"{p}.proj.weight"], sd[f"{p}.proj.bias"])


def hab(x, x_size, sd, p, heads, ws, shift, conv_scale):
    """HAB.forward :266-309. x [b, h*w, c]."""
    h, w = x_size
    b, _, c = x.shape
    shortcut = x
    x = F.layer_norm(x, (c,), sd[f"{p}.norm1.weight"], sd[f"{p}.norm1.bias"], 1e-5).view(b, h, w, c)
    conv_x = cab(x.permute(0, 3, 1, 2), sd, f"{p}.conv_block").permute(0, 2, 3, 1).reshape(b, h * w, c)
    mask = None
    if shift > 0:
        x = torch.roll(x, shifts=(-shift, -shift), dims=(1, 2))
        mask = shift_mask(h, w, ws, shift)
    aw = window_attention(_partition(x, ws), sd, f"{p}.attn", heads, ws, mask)
    x = _reverse(aw, ws, b, h, w)
    if shift > 0:
        x = torch.roll(x, shifts=(shift, shift), dims=(1, 2))
    x = shortcut + x.reshape(b, h * w, c) + conv_x * conv_scale
    return x + _mlp(F.layer_norm(x, (c,), sd[f"{p}.norm2.weight"], sd[f"{p}.norm2.bias"], 1e-5), sd, f"{p}.mlp")


def ocab(x, x_size, sd, p, heads, ws, overlap_ratio):
    """OCAB.forward :392-438: queries from ws x ws windows, keys / values from the (ws + int(ws * ratio))-wide windows
    around them (nn.Unfold, stride ws, ZERO padding applied AFTER the qkv Linear, so padded keys are exact zeros)."""
    h, w = x_size
    b, _, c = x.shape
    ows = int(ws * overlap_ratio) + ws
    d = c // heads
    shortcut = x
    x = F.layer_norm(x, (c,), sd[f"{p}.norm1.weight"], sd[f"{p}.norm1.bias"], 1e-5).view(b, h, w, c)
    qkv = F.linear(x, sd[f"{p}.qkv.weight"], sd[f"{p}.qkv.bias"]).reshape(b, h, w, 3, c).permute(3, 0, 4, 1, 2)
    q = _partition(qkv[0].permute(0, 2, 3, 1), ws)                                 # [nw*b, ws*ws, c]
    kv = torch.cat((qkv[1], qkv[2]), dim=1)                                        # [b, 2c, h, w]
    u = F.unfold(kv, kernel_size=(ows, ows), stride=ws, padding=(ows - ws) // 2)   # [b, 2c*ows*ows, nw]
    nw = u.shape[-1]
    # 'b (nc ch owh oww) nw -> nc (b nw) (owh oww) ch'
    u = u.view(b, 2, c, ows * ows, nw).permute(1, 0, 4, 3, 2).reshape(2, b * nw, ows * ows, c)
    k, v = u[0], u[1]
    b_, nq, n = q.shape[0], ws * ws, ows * ows
    q = q.reshape(b_, nq, heads, d).permute(0, 2, 1, 3) * d ** -0.5
    k = k.reshape(b_, n, heads, d).permute(0, 2, 1, 3)
    v = v.reshape(b_, n, heads, d).permute(0, 2, 1, 3)
    attn = q @ k.transpose(-2, -1)
    bias = sd[f"{p}.relative_position_bias_table"][rpi_oca(ws, ows).reshape(-1)].reshape(nq, n, heads)
    attn = (attn + bias.permute(2, 0, 1).unsqueeze(0)).softmax(-1)
    aw = (attn @ v).transpose(1, 2).reshape(b_, nq, c)
    x = _reverse(aw, ws, b, h, w).reshape(b, h * w, c)
    x = F.linear(x, sd[f"{p}.proj.weight"], sd[f"{p}.proj.bias"]) + shortcut
    return x + _mlp(F.layer_norm(x, (c,), sd[f"{p}.norm2.weight"], sd[f"{p}.norm2.bias"], 1e-5), sd, f"{p}.mlp")


def forward_features(x, sd, cfg):
    """:950-969 with AttenBlocks.forward :526-534 and RHAG.forward :618-619."""
    b, c, h, w = x.shape
    ws = cfg["window_size"]
    t = x.flatten(2).transpose(1, 2)
    t = F.layer_norm(t, (c,), sd["patch_embed.norm.weight"], sd["patch_embed.norm.bias"], 1e-5)
    for r, depth in enumerate(cfg["depths"]):
        res = t
        g = f"layers.{r}.residual_group"
        for i in range(depth):
            t = hab(t, (h, w), sd, f"{g}.blocks.{i}", cfg["num_heads"][r], ws, 0 if i % 2 == 0 else ws // 2,
                    cfg["conv_scale"])
        t = ocab(t, (h, w), sd, f"{g}.overlap_attn", cfg["num_heads"][r], ws, cfg["overlap_ratio"])
        y = F.conv2d(t.transpose(1, 2).reshape(b, c, h, w), sd[f"layers.{r}.conv.weight"], sd[f"layers.{r}.conv.bias"],
                     padding=1)
        t = y.flatten(2).transpose(1, 2) + res
    t = F.layer_norm(t, (c,), sd["norm.weight"], sd["norm.bias"], 1e-5)
    return t.transpose(1, 2).reshape(b, c, h, w)


def forward(sd, cfg, x):
    """HAT.forward :971-984 (upsampler 'pixelshuffle', resi_connection '1conv': the configuration of
    executor/super_resolution/configs/hat.yml). Height and width must be multiples of window_size: the reference
    pads in HATModel.pre_process, not in the network."""
    assert cfg["upsampler"] == "pixelshuffle" and cfg["resi_connection"] == "1conv"
    cin, rng = cfg["in_chans"], cfg["img_range"]
    mean = torch.tensor((0.4488, 0.4371, 0.4040)).view(1, 3, 1, 1) if cin == 3 else torch.zeros(1, 1, 1, 1)
    conv = lambda t, n: F.conv2d(t, sd[f"{n}.weight"], sd[f"{n}.bias"], padding=1)
    x = (x - mean) * rng
    x = conv(x, "conv_first")
    x = conv(forward_features(x, sd, cfg), "conv_after_body") + x
    x = F.leaky_relu(conv(x, "conv_before_upsample.0"), 0.01)
    s = cfg["upscale"]
    assert s & (s - 1) == 0
    for i in range(int(math.log2(s))):
        x = F.pixel_shuffle(conv(x, f"upsample.{2 * i}"), 2)
    x = conv(x, "conv_last")
    return x / rng + mean


class Model:
    """Callable with the signature of the reference nn.Module: model(x[B,c,H,W]) -> [B,c,s*H,s*W]."""

    def __init__(self, sd, cfg):
        self.sd = {k: v.float() for k, v in sd.items()}
        self.cfg = cfg

    @torch.no_grad()
    def __call__(self, x):
        return forward(self.sd, self.cfg, x)
