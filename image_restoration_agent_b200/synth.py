"""Seeded synthetic weights and inputs (no checkpoints ship with the reference and there is no network).

state_dict keys and shapes are exactly those of the reference modules
(restormer_arch.py:193-243; network_swinir.py:618-763 — SURVEY.md §9), so the dicts load with
load_state_dict() into the reference nn.Modules (tools/make_golden.py does that to pin the oracle).
Values are non-trivial on purpose (LayerNorm gains != 1, temperatures != 1, sizeable relative-position biases)
so that parity tests exercise every term.
"""
from __future__ import annotations

import math
from collections import OrderedDict

import numpy as np
import torch


class _Gen:
    def __init__(self, seed: int):
        self.g = torch.Generator(device="cpu")
        self.g.manual_seed(seed)

    def uniform(self, shape, bound):
        return (torch.rand(shape, generator=self.g, dtype=torch.float32) * 2 - 1) * bound

    def normal(self, shape, std, mean=0.0):
        return torch.randn(shape, generator=self.g, dtype=torch.float32) * std + mean

    def conv(self, cout, cin_per_group, k):
        fan_in = cin_per_group * k * k
        return self.uniform((cout, cin_per_group, k, k), 1.0 / math.sqrt(fan_in))

    def linear(self, cout, cin):
        return self.uniform((cout, cin), 1.0 / math.sqrt(cin))


def restormer_state_dict(cfg: dict, seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    g = _Gen(seed)
    sd: OrderedDict[str, torch.Tensor] = OrderedDict()
    dim = cfg["dim"]
    with_bias = cfg["LayerNorm_type"] != "BiasFree"
    assert not cfg.get("bias", False) and not cfg.get("dual_pixel_task", False)

    def block(prefix, C, heads):
        hidden = int(C * cfg["ffn_expansion_factor"])
        for nrm in ("norm1", "norm2"):
            sd[f"{prefix}.{nrm}.body.weight"] = g.normal((C,), 0.1, 1.0)
            if with_bias:
                sd[f"{prefix}.{nrm}.body.bias"] = g.normal((C,), 0.05)
            if nrm == "norm1":
                sd[f"{prefix}.attn.temperature"] = g.uniform((heads, 1, 1), 1.75) + 2.25
                sd[f"{prefix}.attn.qkv.weight"] = g.conv(3 * C, C, 1)
                sd[f"{prefix}.attn.qkv_dwconv.weight"] = g.conv(3 * C, 1, 3)
                sd[f"{prefix}.attn.project_out.weight"] = g.conv(C, C, 1)
        sd[f"{prefix}.ffn.project_in.weight"] = g.conv(2 * hidden, C, 1)
        sd[f"{prefix}.ffn.dwconv.weight"] = g.conv(2 * hidden, 1, 3)
        sd[f"{prefix}.ffn.project_out.weight"] = g.conv(C, hidden, 1)

    nb, heads = cfg["num_blocks"], cfg["heads"]
    sd["patch_embed.proj.weight"] = g.conv(dim, cfg["inp_channels"], 3)
    for i in range(nb[0]):
        block(f"encoder_level1.{i}", dim, heads[0])
    sd["down1_2.body.0.weight"] = g.conv(dim // 2, dim, 3)
    for i in range(nb[1]):
        block(f"encoder_level2.{i}", dim * 2, heads[1])
    sd["down2_3.body.0.weight"] = g.conv(dim, dim * 2, 3)
    for i in range(nb[2]):
        block(f"encoder_level3.{i}", dim * 4, heads[2])
    sd["down3_4.body.0.weight"] = g.conv(dim * 2, dim * 4, 3)
    for i in range(nb[3]):
        block(f"latent.{i}", dim * 8, heads[3])
    sd["up4_3.body.0.weight"] = g.conv(dim * 16, dim * 8, 3)
    sd["reduce_chan_level3.weight"] = g.conv(dim * 4, dim * 8, 1)
    for i in range(nb[2]):
        block(f"decoder_level3.{i}", dim * 4, heads[2])
    sd["up3_2.body.0.weight"] = g.conv(dim * 8, dim * 4, 3)
    sd["reduce_chan_level2.weight"] = g.conv(dim * 2, dim * 4, 1)
    for i in range(nb[1]):
        block(f"decoder_level2.{i}", dim * 2, heads[1])
    sd["up2_1.body.0.weight"] = g.conv(dim * 4, dim * 2, 3)
    for i in range(nb[0]):
        block(f"decoder_level1.{i}", dim * 2, heads[0])
    for i in range(cfg["num_refinement_blocks"]):
        block(f"refinement.{i}", dim * 2, heads[0])
    sd["output.weight"] = g.conv(cfg["out_channels"], dim * 2, 3)
    return sd


def xrestormer_state_dict(cfg: dict, seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    """Keys / shapes of xrestormer_arch.py:324-380 (TransformerBlock :239-260: channel_attn, channel_ffn,
    spatial_attn (OCAB with rel_pos_emb), spatial_ffn, norm1..4)."""
    g = _Gen(seed)
    sd: OrderedDict[str, torch.Tensor] = OrderedDict()
    dim = cfg["dim"]
    with_bias = cfg["LayerNorm_type"] != "BiasFree"
    assert not cfg.get("bias", False) and not cfg.get("dual_pixel_task", False)
    ws, dh = cfg["window_size"], cfg["spatial_dim_head"]

    def ffn(prefix, C):
        hidden = int(C * cfg["ffn_expansion_factor"])
        sd[f"{prefix}.project_in.weight"] = g.conv(2 * hidden, C, 1)
        sd[f"{prefix}.dwconv.weight"] = g.conv(2 * hidden, 1, 3)
        sd[f"{prefix}.project_out.weight"] = g.conv(C, hidden, 1)

    def block(prefix, C, level):
        ch, sh = cfg["channel_heads"][level], cfg["spatial_heads"][level]
        inner = dh * sh
        rel = 2 * (ws + int(ws * cfg["overlap_ratio"][level])) - 1
        # registration order of the reference module (spatial_attn, channel_attn, norm1..4, channel_ffn, spatial_ffn)
        sd[f"{prefix}.spatial_attn.qkv.weight"] = g.conv(3 * inner, C, 1)
        sd[f"{prefix}.spatial_attn.project_out.weight"] = g.conv(C, inner, 1)
        sd[f"{prefix}.spatial_attn.rel_pos_emb.rel_height"] = g.normal((rel, dh), dh ** -0.5)
        sd[f"{prefix}.spatial_attn.rel_pos_emb.rel_width"] = g.normal((rel, dh), dh ** -0.5)
        sd[f"{prefix}.channel_attn.temperature"] = g.uniform((ch, 1, 1), 1.75) + 2.25
        sd[f"{prefix}.channel_attn.qkv.weight"] = g.conv(3 * C, C, 1)
        sd[f"{prefix}.channel_attn.qkv_dwconv.weight"] = g.conv(3 * C, 1, 3)
        sd[f"{prefix}.channel_attn.project_out.weight"] = g.conv(C, C, 1)
        for nrm in ("norm1", "norm2", "norm3", "norm4"):
            sd[f"{prefix}.{nrm}.body.weight"] = g.normal((C,), 0.1, 1.0)
            if with_bias:
                sd[f"{prefix}.{nrm}.body.bias"] = g.normal((C,), 0.05)
        ffn(f"{prefix}.channel_ffn", C)
        ffn(f"{prefix}.spatial_ffn", C)

    nb = cfg["num_blocks"]
    sd["patch_embed.proj.weight"] = g.conv(dim, cfg["inp_channels"], 3)
    for i in range(nb[0]):
        block(f"encoder_level1.{i}", dim, 0)
    sd["down1_2.body.0.weight"] = g.conv(dim // 2, dim, 3)
    for i in range(nb[1]):
        block(f"encoder_level2.{i}", dim * 2, 1)
    sd["down2_3.body.0.weight"] = g.conv(dim, dim * 2, 3)
    for i in range(nb[2]):
        block(f"encoder_level3.{i}", dim * 4, 2)
    sd["down3_4.body.0.weight"] = g.conv(dim * 2, dim * 4, 3)
    for i in range(nb[3]):
        block(f"latent.{i}", dim * 8, 3)
    sd["up4_3.body.0.weight"] = g.conv(dim * 16, dim * 8, 3)
    sd["reduce_chan_level3.weight"] = g.conv(dim * 4, dim * 8, 1)
    for i in range(nb[2]):
        block(f"decoder_level3.{i}", dim * 4, 2)
    sd["up3_2.body.0.weight"] = g.conv(dim * 8, dim * 4, 3)
    sd["reduce_chan_level2.weight"] = g.conv(dim * 2, dim * 4, 1)
    for i in range(nb[1]):
        block(f"decoder_level2.{i}", dim * 2, 1)
    sd["up2_1.body.0.weight"] = g.conv(dim * 4, dim * 2, 3)
    for i in range(nb[0]):
        block(f"decoder_level1.{i}", dim * 2, 0)
    for i in range(cfg["num_refinement_blocks"]):
        block(f"refinement.{i}", dim * 2, 0)
    sd["output.weight"] = g.conv(cfg["out_channels"], dim * 2, 3)
    return sd


def swinir_state_dict(cfg: dict, seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    """Parameters only; the buffers relative_position_index / attn_mask are recomputed analytically."""
    g = _Gen(seed)
    sd: OrderedDict[str, torch.Tensor] = OrderedDict()
    C, ws = cfg["embed_dim"], cfg["window_size"]
    cin = cfg["in_chans"]
    hidden = int(C * cfg["mlp_ratio"])

    def conv(name, cout, cin_, k=3):
        sd[f"{name}.weight"] = g.conv(cout, cin_, k)
        sd[f"{name}.bias"] = g.uniform((cout,), 1.0 / math.sqrt(cin_ * k * k))

    def lin(name, cout, cin_):
        sd[f"{name}.weight"] = g.linear(cout, cin_)
        sd[f"{name}.bias"] = g.uniform((cout,), 1.0 / math.sqrt(cin_))

    def norm(name):
        sd[f"{name}.weight"] = g.normal((C,), 0.1, 1.0)
        sd[f"{name}.bias"] = g.normal((C,), 0.05)

    def resi(name):
        if cfg["resi_connection"] == "1conv":
            conv(name, C, C)
        else:
            conv(f"{name}.0", C // 4, C)
            conv(f"{name}.2", C // 4, C // 4, 1)
            conv(f"{name}.4", C, C // 4)

    if cfg.get("unshuffle"):  # DiffBIR stage 1: Sequential(PixelUnshuffle(sf), Conv2d(cin * sf^2, C)) -> key conv_first.1
        conv("conv_first.1", C, cin * cfg["unshuffle_scale"] ** 2)
    else:
        conv("conv_first", C, cin)
    norm("patch_embed.norm")
    for r, depth in enumerate(cfg["depths"]):
        nh = cfg["num_heads"][r]
        for b in range(depth):
            p = f"layers.{r}.residual_group.blocks.{b}"
            norm(f"{p}.norm1")
            sd[f"{p}.attn.relative_position_bias_table"] = g.normal(((2 * ws - 1) ** 2, nh), 0.5)
            lin(f"{p}.attn.qkv", 3 * C, C)
            lin(f"{p}.attn.proj", C, C)
            norm(f"{p}.norm2")
            lin(f"{p}.mlp.fc1", hidden, C)
            lin(f"{p}.mlp.fc2", C, hidden)
        resi(f"layers.{r}.conv")
    norm("norm")
    resi("conv_after_body")
    up = cfg["upsampler"]
    if up == "pixelshuffle":
        conv("conv_before_upsample.0", 64, C)
        scale = cfg["upscale"]
        if scale & (scale - 1) == 0:
            for i in range(int(math.log2(scale))):
                conv(f"upsample.{2 * i}", 256, 64)
        elif scale == 3:
            conv("upsample.0", 9 * 64, 64)
        else:
            raise ValueError(f"scale {scale} is not supported")
        conv("conv_last", cin, 64)
    elif up == "pixelshuffledirect":
        conv("upsample.0", cfg["upscale"] ** 2 * cin, C)
    elif up == "nearest+conv":
        conv("conv_before_upsample.0", 64, C)
        conv("conv_up1", 64, 64)
        if cfg["upscale"] >= 4:
            conv("conv_up2", 64, 64)
        if cfg["upscale"] == 8:
            conv("conv_up3", 64, 64)
        conv("conv_hr", 64, 64)
        conv("conv_last", cin, 64)
    else:
        conv("conv_last", cin, C)
    return sd


def hat_state_dict(cfg: dict, seed: int = 0) -> "OrderedDict[str, torch.Tensor]":
    """Parameters of hat_arch.HAT (upsampler 'pixelshuffle', '1conv'); the registered buffers
    relative_position_index_SA / _OCA are included because the reference loads with strict=True."""
    g = _Gen(seed)
    sd: OrderedDict[str, torch.Tensor] = OrderedDict()
    C, ws, cin = cfg["embed_dim"], cfg["window_size"], cfg["in_chans"]
    ows = ws + int(cfg["overlap_ratio"] * ws)
    hidden = int(C * cfg["mlp_ratio"])

    def conv(name, cout, cin_, k=3):
        sd[f"{name}.weight"] = g.conv(cout, cin_, k)
        sd[f"{name}.bias"] = g.uniform((cout,), 1.0 / math.sqrt(cin_ * k * k))

    def lin(name, cout, cin_):
        sd[f"{name}.weight"] = g.linear(cout, cin_)
        sd[f"{name}.bias"] = g.uniform((cout,), 1.0 / math.sqrt(cin_))

    def norm(name):
        sd[f"{name}.weight"] = g.normal((C,), 0.1, 1.0)
        sd[f"{name}.bias"] = g.normal((C,), 0.05)

    conv("conv_first", C, cin)
    norm("patch_embed.norm")
    for r, depth in enumerate(cfg["depths"]):
        nh = cfg["num_heads"][r]
        grp = f"layers.{r}.residual_group"
        for b in range(depth):
            p = f"{grp}.blocks.{b}"
            norm(f"{p}.norm1")
            sd[f"{p}.attn.relative_position_bias_table"] = g.normal(((2 * ws - 1) ** 2, nh), 0.5)
            lin(f"{p}.attn.qkv", 3 * C, C)
            lin(f"{p}.attn.proj", C, C)
            conv(f"{p}.conv_block.cab.0", C // cfg["compress_ratio"], C)
            conv(f"{p}.conv_block.cab.2", C, C // cfg["compress_ratio"])
            conv(f"{p}.conv_block.cab.3.attention.1", C // cfg["squeeze_factor"], C, 1)
            conv(f"{p}.conv_block.cab.3.attention.3", C, C // cfg["squeeze_factor"], 1)
            norm(f"{p}.norm2")
            lin(f"{p}.mlp.fc1", hidden, C)
            lin(f"{p}.mlp.fc2", C, hidden)
        p = f"{grp}.overlap_attn"
        norm(f"{p}.norm1")
        lin(f"{p}.qkv", 3 * C, C)
        sd[f"{p}.relative_position_bias_table"] = g.normal(((ws + ows - 1) ** 2, nh), 0.5)
        lin(f"{p}.proj", C, C)
        norm(f"{p}.norm2")
        lin(f"{p}.mlp.fc1", hidden, C)
        lin(f"{p}.mlp.fc2", C, hidden)
        conv(f"layers.{r}.conv", C, C)
    norm("norm")
    conv("conv_after_body", C, C)
    conv("conv_before_upsample.0", 64, C)
    for i in range(int(math.log2(cfg["upscale"]))):
        conv(f"upsample.{2 * i}", 256, 64)
    conv("conv_last", cin, 64)
    return sd


def synthetic_image_u8(h: int, w: int, c: int = 3, seed: int = 0) -> np.ndarray:
    """uint8 HWC image with image-like statistics: uniform noise, 5x5 box-blurred twice, contrast stretched
    (SURVEY.md §8d)."""
    rng = np.random.RandomState(seed)
    x = rng.randint(0, 256, size=(h + 8, w + 8, c)).astype(np.float32)
    for _ in range(2):
        cs = np.cumsum(np.cumsum(np.pad(x, ((1, 0), (1, 0), (0, 0))), axis=0), axis=1)
        x = (cs[5:, 5:] - cs[:-5, 5:] - cs[5:, :-5] + cs[:-5, :-5]) / 25.0
    x = x[:h, :w]
    x = (x - x.min()) / max(float(x.max() - x.min()), 1e-6) * 255.0
    return np.clip(np.rint(x), 0, 255).astype(np.uint8)
