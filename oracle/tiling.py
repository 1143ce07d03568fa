"""ORACLE (test infrastructure, not product code): the reference tools' pad / tile loop / blend / quantise code,
restated with numpy + torch on CPU. `model` is any callable x[B,c,H,W] -> y[B,c,sH,sW] (the oracle models, or the
reference nn.Modules in tools/make_golden.py).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn.functional as F


def tile_indices(n: int, tile: int, overlap: int) -> list[int]:
    """inference.py:305-308 / demo.py:139-141 (tile already min(tile, h, w))."""
    stride = tile - overlap
    return list(range(0, n - tile, stride)) + [n - tile]


def coverage_count(n: int, tile: int, overlap: int, sf: int = 1) -> np.ndarray:
    """The separable integer count that the reference accumulates in W (ones_like per tile)."""
    cnt = np.zeros(n * sf, dtype=np.int64)
    for i in tile_indices(n, tile, overlap):
        cnt[i * sf:(i + tile) * sf] += 1
    return cnt


def tiled_forward(x: torch.Tensor, model, tile, overlap: int, sf: int, multiple: int) -> torch.Tensor:
    """inference.py:294-322 test() / demo.py:129-153. tile=None -> whole image."""
    if tile is None:
        return model(x)
    b, c, h, w = x.shape
    tile = min(tile, h, w)
    assert tile % multiple == 0, "tile size should be a multiple of window_size"
    hs, ws_ = tile_indices(h, tile, overlap), tile_indices(w, tile, overlap)
    E = torch.zeros(b, c, h * sf, w * sf, dtype=x.dtype)
    W = torch.zeros_like(E)
    for hi in hs:
        for wi in ws_:
            out = model(x[..., hi:hi + tile, wi:wi + tile])
            E[..., hi * sf:(hi + tile) * sf, wi * sf:(wi + tile) * sf].add_(out)
            W[..., hi * sf:(hi + tile) * sf, wi * sf:(wi + tile) * sf].add_(torch.ones_like(out))
    return E.div_(W)


def swinir_pad(x: torch.Tensor, ws: int) -> torch.Tensor:
    """inference.py:76-80: symmetric flip-and-concat pad to (h // ws + 1) * ws — always pads."""
    _, _, h, w = x.shape
    hp, wp = (h // ws + 1) * ws - h, (w // ws + 1) * ws - w
    x = torch.cat([x, torch.flip(x, [2])], 2)[:, :, :h + hp, :]
    return torch.cat([x, torch.flip(x, [3])], 3)[:, :, :, :w + wp]


def swinir_process_image(img_bgr_u8: np.ndarray, model, scale: int, window_size: int, tile=None, tile_overlap=32,
                         return_float=False):
    """model_serving.py:131-175 process_image() fed by cv2.imread(...)/255 (inference.py:238-246, 69-90).
    img_bgr_u8: uint8 [H, W, 3] BGR (or [H, W, 1]). Returns uint8 [H*s, W*s, 3] BGR (or [H*s, W*s])."""
    img = img_bgr_u8.astype(np.float32) / 255.0
    img = np.transpose(img if img.shape[2] == 1 else img[:, :, [2, 1, 0]], (2, 0, 1))
    x = torch.from_numpy(np.ascontiguousarray(img)).float().unsqueeze(0)
    _, _, h_old, w_old = x.shape
    x = swinir_pad(x, window_size)
    with torch.no_grad():
        out = tiled_forward(x, model, tile, tile_overlap, scale, window_size)
    out = out[..., :h_old * scale, :w_old * scale]
    if return_float:
        return out[0].clone()
    o = out.squeeze().float().clamp_(0, 1).numpy()
    if o.ndim == 3:
        o = np.transpose(o[[2, 1, 0], :, :], (1, 2, 0))
    return (o * 255.0).round().astype(np.uint8)


def restormer_restore_image(img_rgb_u8: np.ndarray, model, tile=None, tile_overlap=32, return_float=False):
    """demo.py:118-161 / model_serving.py:157-208. img_rgb_u8: uint8 [H, W, c] RGB (or gray c=1).
    Returns uint8 [H, W, c] (img_as_ubyte of the clamped float == rint(x*255))."""
    x = torch.from_numpy(img_rgb_u8).float().div(255.0).permute(2, 0, 1).unsqueeze(0)
    height, width = x.shape[2], x.shape[3]
    m = 8
    H, W = ((height + m) // m) * m, ((width + m) // m) * m
    padh = H - height if height % m != 0 else 0
    padw = W - width if width % m != 0 else 0
    x = F.pad(x, (0, padw, 0, padh), "reflect")
    with torch.no_grad():
        restored = tiled_forward(x, model, tile, tile_overlap, 1, 8)
    if return_float:
        return restored[0, :, :height, :width].clone()
    restored = torch.clamp(restored, 0, 1)[:, :, :height, :width]
    r = restored.permute(0, 2, 3, 1).numpy()[0]
    return np.rint(r * np.float32(255.0)).astype(np.uint8)


def xrestormer_restore_image(img_bgr_u8: np.ndarray, model, return_float=False, padding_size: int = 64, scale: int = 1):
    """The BasicSR test path of the X-Restormer tool (executor/*/configs/xrestormer.yml: no tile, padding_size 64):
    cv2.imread BGR -> RGB float/255 (basicsr img2tensor, bgr2rgb) -> reflect pad bottom/right to a multiple of
    padding_size only when needed (xrestormer_model.py:56-66 pre_process) -> model -> crop (:139-142 post_process) ->
    tensor2img: clamp(0,1), RGB -> BGR, (x * 255).round() -> uint8. img_bgr_u8: uint8 [H, W, 3] BGR; returns BGR."""
    x = torch.from_numpy(np.ascontiguousarray(img_bgr_u8[:, :, ::-1])).float().div(255.0).permute(2, 0, 1).unsqueeze(0)
    _, _, h, w = x.shape
    ph = (padding_size - h % padding_size) % padding_size
    pw = (padding_size - w % padding_size) % padding_size
    x = F.pad(x, (0, pw, 0, ph), "reflect")
    with torch.no_grad():
        y = model(x)
    y = y[:, :, :h * scale, :w * scale]   # post_process: 0 : h - mod_pad_h * scale
    if return_float:
        return y[0].clone()
    o = y[0].float().clamp_(0, 1).numpy()
    o = np.transpose(o[[2, 1, 0], :, :], (1, 2, 0))
    return (o * 255.0).round().astype(np.uint8)


def hat_tile_plan(height: int, width: int, tile_size: int, tile_pad: int) -> list[tuple[int, ...]]:
    """The tile rectangles of HATModel.tile_process (hat/models/hat_model.py:40-75) on the window-padded image, one
    tuple per tile in the reference's y-major order:
    (y0, y1, x0, x1, y0_pad, y1_pad, x0_pad, x1_pad) = the tile's own area and the area fed to the network."""
    plan = []
    for y in range(math.ceil(height / tile_size)):
        for x in range(math.ceil(width / tile_size)):
            x0, y0 = x * tile_size, y * tile_size
            x1, y1 = min(x0 + tile_size, width), min(y0 + tile_size, height)
            plan.append((y0, y1, x0, x1, max(y0 - tile_pad, 0), min(y1 + tile_pad, height), max(x0 - tile_pad, 0),
                         min(x1 + tile_pad, width)))
    return plan


def hat_restore_image(img_bgr_u8: np.ndarray, model, scale: int = 4, window_size: int = 16, tile_size=512, tile_pad=32,
                      return_float=False):
    """The BasicSR test path of the HAT tool (executor/super_resolution/configs/hat.yml): cv2.imread BGR -> RGB
    float/255 (basicsr img2tensor) -> HATModel.pre_process (hat_model.py:16-26: reflect pad bottom/right to a multiple
    of window_size) -> tile_process (:40-105: crop-with-pad tiles, each tile's own area pasted, NO blending;
    tile_size None = process() on the whole image, :28-38) -> post_process (:107-109) -> tensor2img.
    img_bgr_u8: uint8 [H, W, 3] BGR; returns BGR uint8 [H*scale, W*scale, 3]."""
    x = torch.from_numpy(np.ascontiguousarray(img_bgr_u8[:, :, ::-1])).float().div(255.0).permute(2, 0, 1).unsqueeze(0)
    _, c, h, w = x.shape
    ph = (window_size - h % window_size) % window_size
    pw = (window_size - w % window_size) % window_size
    x = F.pad(x, (0, pw, 0, ph), "reflect")
    H, W = x.shape[2:]
    with torch.no_grad():
        if tile_size is None:
            y = model(x)
        else:
            y = x.new_zeros((1, c, H * scale, W * scale))
            for (y0, y1, x0, x1, y0p, y1p, x0p, x1p) in hat_tile_plan(H, W, tile_size, tile_pad):
                o = model(x[:, :, y0p:y1p, x0p:x1p])
                oy, ox = (y0 - y0p) * scale, (x0 - x0p) * scale
                y[:, :, y0 * scale:y1 * scale, x0 * scale:x1 * scale] = o[:, :, oy:oy + (y1 - y0) * scale,
                                                                          ox:ox + (x1 - x0) * scale]
    y = y[:, :, :h * scale, :w * scale]
    if return_float:
        return y[0].clone()
    o = y[0].float().clamp_(0, 1).numpy()
    o = np.transpose(o[[2, 1, 0], :, :], (1, 2, 0))
    return (o * 255.0).round().astype(np.uint8)
