"""ORACLE (test infrastructure, not product code): CPU fp32 restatement of the reference SwinIR forward.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
Restates AgenticIR/executor/denoising/tools/SwinIR/models/network_swinir.py as pure functions over a
state_dict. The relative-position index and the shifted-window mask are computed from token coordinates
(closed forms of :93-103 and :216-237) instead of being read from registered buffers. Pinned against the
reference module itself by tools/make_golden.py -> tests/golden/swinir_*.npz.
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F


def relative_position_index(ws: int) -> torch.Tensor:
    """network_swinir.py:93-103: idx[i,j] = (yi-yj+ws-1)*(2ws-1) + (xi-xj+ws-1), tokens row-major."""
    t = torch.arange(ws * ws)
    y, x = t // ws, t % ws
    return (y[:, None] - y[None, :] + ws - 1) * (2 * ws - 1) + (x[:, None] - x[None, :] + ws - 1)


def shift_mask(H: int, W: int, ws: int, shift: int) -> torch.Tensor:
    """network_swinir.py:216-237: region id 3*rh+rw on the SHIFTED frame, 0 where ids match else -100."""
    def region(n):
        r = torch.zeros(n, dtype=torch.long)
        r[n - ws:n - shift] = 1
        r[n - shift:] = 2
        return r
    ids = (3 * region(H)[:, None] + region(W)[None, :]).float()           # [H, W]
    win = ids.view(H // ws, ws, W // ws, ws).permute(0, 2, 1, 3).reshape(-1, ws * ws)  # window_partition :42-45
    d = win[:, None, :] - win[:, :, None]
    return torch.where(d != 0, torch.full_like(d, -100.0), torch.zeros_like(d))


def window_attention(xw, sd, p, heads, ws, mask):
    """WindowAttention.forward :114-145. xw [nW*B, N, C]."""
    B_, N, C = xw.shape
    qkv = F.linear(xw, sd[f"{p}.qkv.weight"], sd[f"{p}.qkv.bias"]).reshape(B_, N, 3, heads, C // heads)
    q, k, v = qkv.permute(2, 0, 3, 1, 4)
    q = q * (C // heads) ** -0.5
    attn = q @ k.transpose(-2, -1)
    table = sd[f"{p}.relative_position_bias_table"]
    bias = table[relative_position_index(ws).reshape(-1)].reshape(N, N, heads).permute(2, 0, 1)
    attn = attn + bias.unsqueeze(0)
    if mask is not None:
        nW = mask.shape[0]
        attn = attn.view(B_ // nW, nW, heads, N, N) + mask[None, :, None]
        attn = attn.view(-1, heads, N, N)
    attn = attn.softmax(-1)
    out = (attn @ v).transpose(1, 2).reshape(B_, N, C)
    return F.linear(out, sd[f"{p}.proj.weight"], sd[f"{p}.proj.bias"])


def swin_block(x, x_size, sd, p, heads, ws, shift):
    """SwinTransformerBlock.forward :239-279. x [B, L, C]."""
    H, W = x_size
    B, L, C = x.shape
    shortcut = x
    x = F.layer_norm(x, (C,), sd[f"{p}.norm1.weight"], sd[f"{p}.norm1.bias"], 1e-5).view(B, H, W, C)
    if shift > 0:
        x = torch.roll(x, shifts=(-shift, -shift), dims=(1, 2))
    xw = x.view(B, H // ws, ws, W // ws, ws, C).permute(0, 1, 3, 2, 4, 5).reshape(-1, ws * ws, C)
    mask = shift_mask(H, W, ws, shift) if shift > 0 else None
    aw = window_attention(xw, sd, f"{p}.attn", heads, ws, mask)
    x = aw.view(B, H // ws, W // ws, ws, ws, C).permute(0, 1, 3, 2, 4, 5).reshape(B, H, W, C)  # window_reverse
    if shift > 0:
        x = torch.roll(x, shifts=(shift, shift), dims=(1, 2))
    x = shortcut + x.view(B, L, C)
    y = F.layer_norm(x, (C,), sd[f"{p}.norm2.weight"], sd[f"{p}.norm2.bias"], 1e-5)
    y = F.linear(F.gelu(F.linear(y, sd[f"{p}.mlp.fc1.weight"], sd[f"{p}.mlp.fc1.bias"])),
                 sd[f"{p}.mlp.fc2.weight"], sd[f"{p}.mlp.fc2.bias"])  # Mlp :24-30
    return x + y


def _resi_conv(x, sd, name, three):
    """'1conv' / '3conv' residual convolution (:464-471, :728-736)."""
    if not three:
        return F.conv2d(x, sd[f"{name}.weight"], sd[f"{name}.bias"], padding=1)
    x = F.leaky_relu(F.conv2d(x, sd[f"{name}.0.weight"], sd[f"{name}.0.bias"], padding=1), 0.2)
    x = F.leaky_relu(F.conv2d(x, sd[f"{name}.2.weight"], sd[f"{name}.2.bias"]), 0.2)
    return F.conv2d(x, sd[f"{name}.4.weight"], sd[f"{name}.4.bias"], padding=1)


def effective_window(cfg):
    """SwinTransformerBlock.__init__ :193-196 uses the CONSTRUCTOR img_size: if it is <= window_size the window
    shrinks to it and the shift is disabled."""
    ws = cfg["window_size"]
    if cfg["img_size"] <= ws:
        return cfg["img_size"], False
    return ws, True


def forward_features(x, sd, cfg):
    """:790-803. x [B, C, H, W] -> [B, C, H, W]."""
    B, C, H, W = x.shape
    ws, can_shift = effective_window(cfg)
    three = cfg["resi_connection"] == "3conv"
    t = x.flatten(2).transpose(1, 2)
    t = F.layer_norm(t, (C,), sd["patch_embed.norm.weight"], sd["patch_embed.norm.bias"], 1e-5)
    for r, depth in enumerate(cfg["depths"]):
        res = t
        for b in range(depth):
            shift = (ws // 2) if (b % 2 == 1 and can_shift) else 0   # :383
            t = swin_block(t, (H, W), sd, f"layers.{r}.residual_group.blocks.{b}", cfg["num_heads"][r], ws, shift)
        y = _resi_conv(t.transpose(1, 2).reshape(B, C, H, W), sd, f"layers.{r}.conv", three)  # RSTB :481-482
        t = y.flatten(2).transpose(1, 2) + res
    t = F.layer_norm(t, (C,), sd["norm.weight"], sd["norm.bias"], 1e-5)
    return t.transpose(1, 2).reshape(B, C, H, W)


def forward(sd, cfg, x):
    """SwinIR.forward :805-840."""
    H, W = x.shape[2:]
    ws = cfg["window_size"]
    x = F.pad(x, (0, (ws - W % ws) % ws, 0, (ws - H % ws) % ws), "reflect")  # check_image_size :783-788
    cin, rng, up = cfg["in_chans"], cfg["img_range"], cfg["upsampler"]
    mean = torch.tensor((0.4488, 0.4371, 0.4040)).view(1, 3, 1, 1) if cin == 3 else torch.zeros(1, 1, 1, 1)
    x = (x - mean) * rng
    three = cfg["resi_connection"] == "3conv"
    conv = lambda t, n: F.conv2d(t, sd[f"{n}.weight"], sd[f"{n}.bias"], padding=1)
    body = lambda f: _resi_conv(forward_features(f, sd, cfg), sd, "conv_after_body", three) + f
    if up == "pixelshuffle":
        x = body(conv(x, "conv_first"))
        x = F.leaky_relu(conv(x, "conv_before_upsample.0"), 0.01)  # nn.LeakyReLU() default slope :742-743
        s = cfg["upscale"]
        if s & (s - 1) == 0:
            for i in range(int(math.log2(s))):
                x = F.pixel_shuffle(conv(x, f"upsample.{2 * i}"), 2)
        else:
            x = F.pixel_shuffle(conv(x, "upsample.0"), 3)
        x = conv(x, "conv_last")
    elif up == "pixelshuffledirect":
        x = body(conv(x, "conv_first"))
        x = F.pixel_shuffle(conv(x, "upsample.0"), cfg["upscale"])
    elif up == "nearest+conv":
        if cfg.get("unshuffle"):
            # DiffBIR's stage-1 SwinIR (DiffBIR/model/swinir.py:716-724, :868-893): conv_first = Sequential(
            # PixelUnshuffle(sf), Conv2d(in_chans * sf^2, C)); the tail up-samples by the same sf (three nearest x2 stages)
            x = body(conv(F.pixel_unshuffle(x, cfg["unshuffle_scale"]), "conv_first.1"))
        else:
            x = body(conv(x, "conv_first"))
        x = F.leaky_relu(conv(x, "conv_before_upsample.0"), 0.01)
        x = F.leaky_relu(conv(F.interpolate(x, scale_factor=2, mode="nearest"), "conv_up1"), 0.2)
        if cfg["upscale"] >= 4:
            x = F.leaky_relu(conv(F.interpolate(x, scale_factor=2, mode="nearest"), "conv_up2"), 0.2)
        if cfg["upscale"] == 8:
            x = F.leaky_relu(conv(F.interpolate(x, scale_factor=2, mode="nearest"), "conv_up3"), 0.2)
        x = conv(F.leaky_relu(conv(x, "conv_hr"), 0.2), "conv_last")
    else:
        x_first = conv(x, "conv_first")
        x = x + conv(body(x_first), "conv_last")
    x = x / rng + mean
    return x[:, :, :H * cfg["upscale"], :W * cfg["upscale"]]  # DiffBIR: H * 8 exceeds the tensor, i.e. no crop (:895)


class Model:
    """Callable with the signature of the reference nn.Module."""

    def __init__(self, sd, cfg):
        self.sd = {k: v.float() for k, v in sd.items()}
        self.cfg = cfg

    @torch.no_grad()
    def __call__(self, x):
        return forward(self.sd, self.cfg, x)
