"""Resident model handle over the C ABI: build once, restore many images.

torch is used for device memory and streams only; all arithmetic happens inside libirb200.so.
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np
import torch

from . import _lib
from ._lib import ARCH_RESTORMER, ARCH_SWINIR, BF16, F32, ModelDesc, check
from .weights import state_dict_to_blob

_UPSAMPLER = {"": 0, None: 0, "pixelshuffle": 1, "nearest+conv": 2}


def restormer_desc(cfg: dict, precision: int) -> ModelDesc:
    if cfg.get("bias", False) or cfg.get("dual_pixel_task", False):
        raise ValueError("irb200 Restormer supports bias=False, dual_pixel_task=False (all agent tasks, demo.py:92)")
    if cfg["inp_channels"] != cfg["out_channels"]:
        raise ValueError("inp_channels must equal out_channels")
    d = ModelDesc()
    d.arch, d.precision, d.in_chans = ARCH_RESTORMER, precision, cfg["inp_channels"]
    d.dim = cfg["dim"]
    d.num_blocks[:] = cfg["num_blocks"]
    d.num_refinement_blocks = cfg["num_refinement_blocks"]
    d.heads[:] = cfg["heads"]
    d.ffn_expansion_factor = cfg["ffn_expansion_factor"]
    d.ln_bias_free = int(cfg["LayerNorm_type"] == "BiasFree")
    return d


def swinir_desc(cfg: dict, precision: int) -> ModelDesc:
    if cfg["upsampler"] not in _UPSAMPLER:
        raise ValueError(f"upsampler {cfg['upsampler']!r} is not served by irb200 (agent tools use '', "
                         "'pixelshuffle', 'nearest+conv')")
    d = ModelDesc()
    d.arch, d.precision, d.in_chans = ARCH_SWINIR, precision, cfg["in_chans"]
    d.img_size, d.window_size, d.embed_dim = cfg["img_size"], cfg["window_size"], cfg["embed_dim"]
    d.num_layers = len(cfg["depths"])
    for i, (dep, nh) in enumerate(zip(cfg["depths"], cfg["num_heads"])):
        d.depths[i], d.num_heads[i] = dep, nh
    d.mlp_ratio = int(cfg["mlp_ratio"])
    d.upscale = cfg["upscale"]
    d.img_range = cfg["img_range"]
    d.upsampler = _UPSAMPLER[cfg["upsampler"]]
    d.resi_3conv = int(cfg["resi_connection"] == "3conv")
    return d


class Engine:
    """One model resident on one GPU. Thread-safe: concurrent Flask requests serialise on a lock
    (the reference shares one model object across threads with no lock, model_serving.py:39,105-128)."""

    def __init__(self, arch: str, cfg: dict, state_dict, precision: str = "bf16", device: int = 0):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise _lib.IrbError("irb200 needs an sm_100 GPU: there is no CPU fallback")
        self.arch, self.cfg, self.device = arch, cfg, device
        prec = {"bf16": BF16, "fp32": F32, "f32": F32}[precision]
        self.precision = precision
        desc = restormer_desc(cfg, prec) if arch == "restormer" else swinir_desc(cfg, prec)
        blob = state_dict_to_blob(state_dict)
        h = C.c_void_p()
        torch.cuda.set_device(device)
        check(self.lib.irb_create(C.byref(desc), blob, len(blob), device, C.byref(h)))
        self._h = h
        self.scale = self.lib.irb_scale(h)
        self.channels = desc.in_chans
        self._ws = None
        self._lock = threading.Lock()

    def close(self):
        if getattr(self, "_h", None):
            self.lib.irb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def set_tile_batch(self, n: int):
        check(self.lib.irb_set_option(self._h, b"tile_batch", int(n)))

    def _workspace(self, nbytes: int) -> torch.Tensor:
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = None
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=f"cuda:{self.device}")
        return self._ws

    # ---- model(x): the nn.Module call of the reference (network_swinir.py:805, restormer_arch.py:245) ----
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        assert x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.shape[1] == self.channels
        x = x.contiguous()
        n, c, h, w = x.shape
        y = torch.empty((n, c, h * self.scale, w * self.scale), device=x.device, dtype=torch.float32)
        need = C.c_size_t()
        with self._lock:
            check(self.lib.irb_workspace_bytes(self._h, n, h, w, C.byref(need)))
            ws = self._workspace(need.value)
            check(self.lib.irb_forward_tiles(self._h, x.data_ptr(), y.data_ptr(), n, h, w, ws.data_ptr(), ws.numel(),
                                             torch.cuda.current_stream().cuda_stream))
        return y

    __call__ = forward

    # ---- the whole tool body for one image ----
    def restore_device(self, img: torch.Tensor, tile=None, tile_overlap: int = 32, bgr: bool = False,
                       want_u8: bool = True, want_f32: bool = False):
        """img: CUDA uint8 [H,W,c] or CUDA fp32 planar [c,H,W] in [0,1]. Returns (u8 [sH,sW,c] | None, f32 | None)."""
        if img.dtype == torch.uint8:
            h, w, c = img.shape
            fmt = 0
        else:
            c, h, w = img.shape
            fmt = 1
        assert c == self.channels, f"model has {self.channels} channel(s), image has {c}"
        img = img.contiguous()
        s = self.scale
        u8 = torch.empty((h * s, w * s, c), dtype=torch.uint8, device=img.device) if want_u8 else None
        f32 = torch.empty((c, h * s, w * s), dtype=torch.float32, device=img.device) if want_f32 else None
        t = 0 if tile is None else int(tile)
        need = C.c_size_t()
        with self._lock:
            check(self.lib.irb_restore_workspace_bytes(self._h, h, w, t, tile_overlap, C.byref(need)))
            ws = self._workspace(need.value)
            check(self.lib.irb_restore_image(self._h, img.data_ptr(), fmt, int(bgr), h, w, t, tile_overlap,
                                             None if u8 is None else u8.data_ptr(),
                                             None if f32 is None else f32.data_ptr(), ws.data_ptr(), ws.numel(),
                                             torch.cuda.current_stream().cuda_stream))
        return u8, f32

    # ---- the stages of restore_device, for sharding tiles across ranks (parallel.py) ----
    def plan(self, h: int, w: int, tile=None, tile_overlap: int = 32) -> dict:
        ys, xs = (C.c_int * 128)(), (C.c_int * 128)()
        v = [C.c_int() for _ in range(6)]
        hp, wp, th, tw, ny, nx = v
        check(self.lib.irb_restore_plan(self._h, h, w, 0 if tile is None else int(tile), tile_overlap, C.byref(hp),
                                        C.byref(wp), C.byref(th), C.byref(tw), ys, C.byref(ny), xs, C.byref(nx)))
        return dict(hp=hp.value, wp=wp.value, th=th.value, tw=tw.value, ys=list(ys[:ny.value]),
                    xs=list(xs[:nx.value]), scale=self.scale, c=self.channels, h=h, w=w)

    def prepare_tiles(self, img: torch.Tensor, plan: dict, tile=None, tile_overlap: int = 32, bgr: bool = False):
        """-> fp32 channels-last tiles [ny*nx, th, tw, c] of the padded image (pad rule of the reference tool)."""
        fmt = 0 if img.dtype == torch.uint8 else 1
        img = img.contiguous()
        c = self.channels
        padded = torch.empty((c, plan["hp"], plan["wp"]), dtype=torch.float32, device=img.device)
        nt = len(plan["ys"]) * len(plan["xs"])
        tiles = torch.empty((nt, plan["th"], plan["tw"], c), dtype=torch.float32, device=img.device)
        check(self.lib.irb_tiles_prepare(self._h, img.data_ptr(), fmt, int(bgr), plan["h"], plan["w"],
                                         0 if tile is None else int(tile), tile_overlap, padded.data_ptr(),
                                         tiles.data_ptr(), torch.cuda.current_stream().cuda_stream))
        return tiles

    def forward_cl(self, tiles: torch.Tensor) -> torch.Tensor:
        """model(x) on channels-last fp32 tiles [n, th, tw, c] -> [n, th*s, tw*s, c]."""
        n, th, tw, c = tiles.shape
        s = self.scale
        out = torch.empty((n, th * s, tw * s, c), dtype=torch.float32, device=tiles.device)
        if n == 0:
            return out
        need = C.c_size_t()
        with self._lock:
            check(self.lib.irb_workspace_bytes(self._h, n, th, tw, C.byref(need)))
            ws = self._workspace(need.value)
            check(self.lib.irb_forward_tiles_cl(self._h, tiles.data_ptr(), out.data_ptr(), n, th, tw, ws.data_ptr(),
                                                ws.numel(), torch.cuda.current_stream().cuda_stream))
        return out

    def stitch(self, tiles_out: torch.Tensor, plan: dict, bgr: bool = False, want_u8=True, want_f32=False):
        from . import ops
        s = plan["scale"]
        f32, u8 = ops.stitch(tiles_out, plan["ys"], plan["xs"], s, plan["hp"], plan["wp"], plan["h"] * s,
                             plan["w"] * s, want_u8=want_u8, reverse_channels=bgr)
        return u8, (f32 if want_f32 else None)

    def restore(self, img_u8: np.ndarray, tile=None, tile_overlap: int = 32, bgr: bool = False) -> np.ndarray:
        """Host uint8 [H,W,c] (or [H,W]) in -> host uint8 out; H2D and D2H included."""
        squeeze = img_u8.ndim == 2
        if squeeze:
            img_u8 = img_u8[:, :, None]
        d = torch.from_numpy(np.ascontiguousarray(img_u8)).to(f"cuda:{self.device}", non_blocking=True)
        u8, _ = self.restore_device(d, tile, tile_overlap, bgr)
        out = u8.cpu().numpy()
        return out[:, :, 0] if squeeze else out
