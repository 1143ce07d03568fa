"""Model configurations of the reference tools, as plain dicts (same keys as the reference constructors).

Restormer: AgenticIR/executor/defocus_deblurring/tools/Restormer/demo.py:52-70, 92
SwinIR   : AgenticIR/executor/denoising/tools/SwinIR/inference.py:136-195 (define_model) and
           AgentApp/model_service/denoising/SwinIR/model_serving.py:43-102
"""
from __future__ import annotations

import copy

RESTORMER_BASE = {
    "inp_channels": 3, "out_channels": 3, "dim": 48, "num_blocks": [4, 6, 6, 8], "num_refinement_blocks": 4,
    "heads": [1, 2, 4, 8], "ffn_expansion_factor": 2.66, "bias": False, "LayerNorm_type": "WithBias",
    "dual_pixel_task": False,
}

RESTORMER_TASKS = ("Motion_Deblurring", "Single_Image_Defocus_Deblurring", "Deraining", "Real_Denoising",
                   "Gaussian_Gray_Denoising", "Gaussian_Color_Denoising")


def restormer_config(task: str) -> dict:
    """demo.py:52-70 get_weights_and_parameters()."""
    if task not in RESTORMER_TASKS:
        raise ValueError(f"Not supported task {task}")
    p = copy.deepcopy(RESTORMER_BASE)
    if task in ("Real_Denoising", "Gaussian_Color_Denoising", "Gaussian_Gray_Denoising"):
        p["LayerNorm_type"] = "BiasFree"
    if task == "Gaussian_Gray_Denoising":
        p["inp_channels"] = 1
        p["out_channels"] = 1
    return p


XRESTORMER_BASE = {  # executor/{dehazing,denoising,deraining,motion_deblurring}/configs/xrestormer.yml: network_g
    "inp_channels": 3, "out_channels": 3, "dim": 48, "num_blocks": [2, 4, 4, 4], "num_refinement_blocks": 4,
    "channel_heads": [1, 2, 4, 8], "spatial_heads": [1, 2, 4, 8], "overlap_ratio": [0.5, 0.5, 0.5, 0.5],
    "window_size": 8, "spatial_dim_head": 16, "ffn_expansion_factor": 2.66, "bias": False,
    "LayerNorm_type": "WithBias", "dual_pixel_task": False, "scale": 1,
}
XRESTORMER_SUBTASKS = ("dehazing", "denoising", "deraining", "motion_deblurring", "super_resolution")


def xrestormer_config(subtask: str = "dehazing") -> dict:
    """The one architecture all X-Restormer tools share; scale 1 / padding_size 64 in four ymls, scale 4 / padding_size
    16 in executor/super_resolution/configs/xrestormer.yml (the input is bilinearly up-sampled x4 first)."""
    p = copy.deepcopy(XRESTORMER_BASE)
    if subtask == "super_resolution":
        p["scale"] = 4
    return p


def small_xrestormer_config(ln_type="WithBias", scale=1) -> dict:
    p = copy.deepcopy(XRESTORMER_BASE)
    p.update(num_blocks=[1, 1, 1, 1], num_refinement_blocks=1, LayerNorm_type=ln_type, scale=scale)
    return p


SWINIR_TASKS = ("classical_sr", "lightweight_sr", "real_sr", "gray_dn", "color_dn", "jpeg_car", "color_jpeg_car")


def swinir_config(task: str, scale: int = 1, large_model: bool = False, training_patch_size: int = 48) -> dict:
    """inference.py:136-195 define_model(): constructor kwargs of SwinIR for each task."""
    m6 = dict(depths=[6] * 6, embed_dim=180, num_heads=[6] * 6, mlp_ratio=2, resi_connection="1conv")
    if task == "classical_sr":
        return dict(upscale=scale, in_chans=3, img_size=training_patch_size, window_size=8, img_range=1.0,
                    upsampler="pixelshuffle", **m6)
    if task == "lightweight_sr":
        return dict(upscale=scale, in_chans=3, img_size=64, window_size=8, img_range=1.0, depths=[6] * 4,
                    embed_dim=60, num_heads=[6] * 4, mlp_ratio=2, upsampler="pixelshuffledirect",
                    resi_connection="1conv")
    if task == "real_sr":
        if not large_model:
            return dict(upscale=scale, in_chans=3, img_size=64, window_size=8, img_range=1.0,
                        upsampler="nearest+conv", **m6)
        return dict(upscale=scale, in_chans=3, img_size=64, window_size=8, img_range=1.0, depths=[6] * 9,
                    embed_dim=240, num_heads=[8] * 9, mlp_ratio=2, upsampler="nearest+conv",
                    resi_connection="3conv")
    if task == "gray_dn":
        return dict(upscale=1, in_chans=1, img_size=128, window_size=8, img_range=1.0, upsampler="", **m6)
    if task == "color_dn":
        return dict(upscale=1, in_chans=3, img_size=128, window_size=8, img_range=1.0, upsampler="", **m6)
    if task == "jpeg_car":
        return dict(upscale=1, in_chans=1, img_size=126, window_size=7, img_range=255.0, upsampler="", **m6)
    if task == "color_jpeg_car":
        return dict(upscale=1, in_chans=3, img_size=126, window_size=7, img_range=255.0, upsampler="", **m6)
    raise ValueError(f"Unknown task: {task}")


def diffbir_swinir_config() -> dict:
    """DiffBIR's stage-1 degradation-removal SwinIR (executor/super_resolution/configs/diffbir.yml:92-107,
    DiffBIR/model/swinir.py:629-800): PixelUnshuffle(8) head, 8 RSTB, nearest+conv x8 tail; output size == input size."""
    return dict(upscale=8, in_chans=3, img_size=64, window_size=8, img_range=1.0, depths=[6] * 8, embed_dim=180,
                num_heads=[6] * 8, mlp_ratio=2, upsampler="nearest+conv", resi_connection="1conv", unshuffle=True,
                unshuffle_scale=8)


def small_diffbir_swinir_config() -> dict:
    p = diffbir_swinir_config()
    p["depths"], p["num_heads"] = [2, 2], [6, 6]
    return p


def hat_config() -> dict:
    """HAT (Real_HAT_GAN_SRx4): executor/super_resolution/configs/hat.yml:21-37 (network_g) with the tile settings of
    :7-9 (tile_size 512, tile_pad 32) under "tile"."""
    return dict(upscale=4, in_chans=3, img_size=64, window_size=16, compress_ratio=3, squeeze_factor=30, conv_scale=0.01,
                overlap_ratio=0.5, img_range=1.0, depths=[6] * 6, embed_dim=180, num_heads=[6] * 6, mlp_ratio=2,
                upsampler="pixelshuffle", resi_connection="1conv")


HAT_TILE = dict(tile_size=512, tile_pad=32)  # hat.yml:7-9


def small_hat_config(window_size=16, upscale=4) -> dict:
    """Shallow HAT of the same structure (2 RHAGs x 2 HABs + OCAB, C = 60, 6 heads -> head_dim 10) for CPU-speed tests."""
    return dict(upscale=upscale, in_chans=3, img_size=64, window_size=window_size, compress_ratio=3, squeeze_factor=30,
                conv_scale=0.01, overlap_ratio=0.5, img_range=1.0, depths=[2, 2], embed_dim=60, num_heads=[6, 6],
                mlp_ratio=2, upsampler="pixelshuffle", resi_connection="1conv")


def small_restormer_config(ln_type="WithBias", inp_channels=3) -> dict:
    """A shrunken Restormer (same code paths, every level present) for fast parity tests."""
    p = copy.deepcopy(RESTORMER_BASE)
    p.update(num_blocks=[1, 1, 1, 2], num_refinement_blocks=1, LayerNorm_type=ln_type, inp_channels=inp_channels,
             out_channels=inp_channels)
    return p


def small_swinir_config(task="color_dn", scale=1, large_model=False) -> dict:
    """A shrunken SwinIR (2 RSTB x 2 STL, so shifted and unshifted blocks both run) for fast parity tests."""
    p = swinir_config(task, scale=scale, large_model=large_model)
    n = 2
    p["depths"] = [2] * n
    p["num_heads"] = p["num_heads"][:n]
    return p
