"""Resident model handle over the C ABI: build once, restore many images.

torch is used for device memory and streams only; all arithmetic happens inside libirb200.so.
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np
import torch

from . import _lib
from ._lib import ARCH_HAT, ARCH_RESTORMER, ARCH_SWINIR, ARCH_XRESTORMER, BF16, F32, ModelDesc, check
from .weights import state_dict_to_blob

_UPSAMPLER = {"": 0, None: 0, "pixelshuffle": 1, "nearest+conv": 2, "pixelshuffledirect": 3}


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


def xrestormer_desc(cfg: dict, precision: int) -> ModelDesc:
    """XRestormer(...) of xrestormer_arch.py:324-341 with the agent's configs (executor/*/configs/xrestormer.yml)."""
    if cfg.get("bias", False) or cfg.get("dual_pixel_task", False) or cfg.get("scale", 1) not in (1, 2, 4):
        raise ValueError("irb200 X-Restormer supports bias=False, dual_pixel_task=False, scale in (1, 2, 4)")
    if len(set(cfg["overlap_ratio"])) != 1:
        raise ValueError("irb200 X-Restormer needs one overlap_ratio for all levels")
    d = ModelDesc()
    d.arch, d.precision, d.in_chans = ARCH_XRESTORMER, precision, cfg["inp_channels"]
    d.dim = cfg["dim"]
    d.num_blocks[:] = cfg["num_blocks"]
    d.num_refinement_blocks = cfg["num_refinement_blocks"]
    d.heads[:] = cfg["channel_heads"]
    d.ffn_expansion_factor = cfg["ffn_expansion_factor"]
    d.ln_bias_free = int(cfg["LayerNorm_type"] == "BiasFree")
    d.window_size = cfg["window_size"]
    d.spatial_heads[:] = cfg["spatial_heads"]
    d.spatial_dim_head = cfg["spatial_dim_head"]
    d.overlap_ratio = cfg["overlap_ratio"][0]
    d.upscale = cfg.get("scale", 1)
    return d


def swinir_desc(cfg: dict, precision: int) -> ModelDesc:
    if cfg["upsampler"] not in _UPSAMPLER:
        raise ValueError(f"unknown SwinIR upsampler {cfg['upsampler']!r}")
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
    d.unshuffle_scale = int(cfg["unshuffle_scale"]) if cfg.get("unshuffle") else 0
    return d


def hat_desc(cfg: dict, precision: int) -> ModelDesc:
    """HAT(...) of hat_arch.py:738-766 with executor/super_resolution/configs/hat.yml:21-37."""
    if cfg["upsampler"] != "pixelshuffle" or cfg["resi_connection"] != "1conv":
        raise ValueError("irb200 serves HAT with upsampler 'pixelshuffle' and resi_connection '1conv' (hat.yml)")
    d = swinir_desc(cfg, precision)
    d.arch = ARCH_HAT
    d.overlap_ratio = float(cfg["overlap_ratio"])
    d.compress_ratio, d.squeeze_factor = int(cfg["compress_ratio"]), int(cfg["squeeze_factor"])
    d.conv_scale = float(cfg["conv_scale"])
    return d


class Engine:
    """One model resident on one GPU. Thread-safe: concurrent Flask requests serialise on a lock
    (the reference shares one model object across threads with no lock, model_serving.py:39,105-128)."""

    MAX_GRAPHS = 8  # captured (shape, tile, overlap, flags) combinations kept per engine, oldest dropped first

    def __init__(self, arch: str, cfg: dict, state_dict, precision: str = "bf16", device: int = 0):
        self.lib = _lib.load()
        if not torch.cuda.is_available():
            raise _lib.IrbError("irb200 needs an sm_100 GPU: there is no CPU fallback")
        self.arch, self.cfg, self.device = arch, cfg, int(device)
        self.dev = torch.device("cuda", self.device)
        prec = {"bf16": BF16, "fp32": F32, "f32": F32}[precision]
        self.precision = precision
        desc = {"restormer": restormer_desc, "swinir": swinir_desc, "xrestormer": xrestormer_desc,
                "hat": hat_desc}[arch](cfg, prec)
        blob = state_dict_to_blob(state_dict)
        h = C.c_void_p()
        # irb_create uploads on `device` and restores the caller's current device; the calling thread's device is
        # never changed by this class (every call below runs under torch.cuda.device(self.device) and the library
        # entry points switch to the model's device themselves).
        check(self.lib.irb_create(C.byref(desc), blob, len(blob), self.device, C.byref(h)))
        self._h = h
        self.scale = self.lib.irb_scale(h)
        self.channels = desc.in_chans
        self.strict_tile = True   # default of restore_device(strict_tile=None); False = process_image semantics
        self._ws = None
        self._ws_gen = 0          # bumped when the workspace is re-allocated: captured graphs hold its address
        self._graphs: dict = {}
        self._lock = threading.Lock()

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.dev).cuda_stream

    def _on_device(self, *tensors):
        for t in tensors:
            if t is not None and t.device != self.dev:
                raise ValueError(f"irb200: tensor on {t.device}, engine on {self.dev}")

    def close(self):
        if getattr(self, "_graphs", None):
            self._graphs.clear()
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
            self._graphs.clear()
            self._ws_gen += 1
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=self.dev)
        return self._ws

    # ---- model(x): the nn.Module call of the reference (network_swinir.py:805, restormer_arch.py:245) ----
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        assert x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.shape[1] == self.channels
        self._on_device(x)
        x = x.contiguous()
        n, c, h, w = x.shape
        y = torch.empty((n, c, h * self.scale, w * self.scale), device=x.device, dtype=torch.float32)
        need = C.c_size_t()
        with self._lock, torch.cuda.device(self.dev):
            check(self.lib.irb_workspace_bytes(self._h, n, h, w, C.byref(need)))
            ws = self._workspace(need.value)
            check(self.lib.irb_forward_tiles(self._h, x.data_ptr(), y.data_ptr(), n, h, w, ws.data_ptr(), ws.numel(),
                                             self._stream()))
        return y

    __call__ = forward

    # ---- the whole tool body for one image ----
    def _restore_call(self, img, fmt, bgr, h, w, t, tile_overlap, u8, f32, strict=None):
        need = C.c_size_t()
        check(self.lib.irb_set_option(self._h, b"strict_tile", int(self.strict_tile if strict is None else strict)))
        check(self.lib.irb_restore_workspace_bytes(self._h, h, w, t, tile_overlap, C.byref(need)))
        ws = self._workspace(need.value)
        check(self.lib.irb_restore_image(self._h, img.data_ptr(), fmt, int(bgr), h, w, t, tile_overlap,
                                         None if u8 is None else u8.data_ptr(),
                                         None if f32 is None else f32.data_ptr(), ws.data_ptr(), ws.numel(),
                                         self._stream()))

    def restore_device(self, img: torch.Tensor, tile=None, tile_overlap: int = 32, bgr: bool = False,
                       want_u8: bool = True, want_f32: bool = False, graph: bool = False, strict_tile=None):
        """img: CUDA uint8 [H,W,c] or CUDA fp32 planar [c,H,W] in [0,1]. Returns (u8 [sH,sW,c] | None, f32 | None).

        graph=True replays a CUDA graph captured once per (image shape, tile, overlap, flags): the ~300-450 kernel
        launches of a forward are then issued by one cudaGraphLaunch. The returned tensors are the
        graph's static output buffers: they are overwritten by the next graph=True call with the same key."""
        if img.dtype == torch.uint8:
            h, w, c = img.shape
            fmt = 0
        else:
            c, h, w = img.shape
            fmt = 1
        assert c == self.channels, f"model has {self.channels} channel(s), image has {c}"
        self._on_device(img)
        img = img.contiguous()
        s = self.scale
        t = 0 if tile is None else int(tile)
        with self._lock, torch.cuda.device(self.dev):
            if not graph:
                u8 = torch.empty((h * s, w * s, c), dtype=torch.uint8, device=self.dev) if want_u8 else None
                f32 = torch.empty((c, h * s, w * s), dtype=torch.float32, device=self.dev) if want_f32 else None
                self._restore_call(img, fmt, bgr, h, w, t, tile_overlap, u8, f32, strict_tile)
                return u8, f32
            key = (fmt, h, w, t, int(tile_overlap), bool(bgr), bool(want_u8), bool(want_f32),
                   bool(self.strict_tile if strict_tile is None else strict_tile))
            g = self._graphs.get(key)
            if g is None or g["gen"] != self._ws_gen:
                g = self._capture(key, img)
            g["in"].copy_(img, non_blocking=True)
            g["graph"].replay()
            return g["u8"], g["f32"]

    def restore_crop_device(self, img: torch.Tensor, tile_size=None, tile_pad: int = 32, bgr: bool = False,
                            want_u8: bool = True, want_f32: bool = False):
        """HATModel's tool body (hat_model.py:16-109): window padding, crop-with-pad tiles (tile_size None = the whole
        image), each tile's own area pasted without blending, crop. Tensors as in restore_device."""
        if img.dtype == torch.uint8:
            h, w, c = img.shape
            fmt = 0
        else:
            c, h, w = img.shape
            fmt = 1
        assert c == self.channels, f"model has {self.channels} channel(s), image has {c}"
        self._on_device(img)
        img = img.contiguous()
        s = self.scale
        t = 0 if tile_size is None else int(tile_size)
        need = C.c_size_t()
        with self._lock, torch.cuda.device(self.dev):
            u8 = torch.empty((h * s, w * s, c), dtype=torch.uint8, device=self.dev) if want_u8 else None
            f32 = torch.empty((c, h * s, w * s), dtype=torch.float32, device=self.dev) if want_f32 else None
            check(self.lib.irb_restore_crop_workspace_bytes(self._h, h, w, t, int(tile_pad), C.byref(need)))
            ws = self._workspace(need.value)
            check(self.lib.irb_restore_image_crop(self._h, img.data_ptr(), fmt, int(bgr), h, w, t, int(tile_pad),
                                                  None if u8 is None else u8.data_ptr(),
                                                  None if f32 is None else f32.data_ptr(), ws.data_ptr(), ws.numel(),
                                                  self._stream()))
        return u8, f32

    def restore_crop(self, img_u8: np.ndarray, tile_size=None, tile_pad: int = 32, bgr: bool = False) -> np.ndarray:
        """Host uint8 [H,W,c] in -> host uint8 [sH,sW,c] out through restore_crop_device; H2D and D2H included."""
        d = torch.from_numpy(np.ascontiguousarray(img_u8)).to(self.dev, non_blocking=True)
        u8, _ = self.restore_crop_device(d, tile_size, tile_pad, bgr)
        return u8.cpu().numpy()

    def _capture(self, key, img):
        fmt, h, w, t, ov, bgr, want_u8, want_f32, strict = key
        s, c = self.scale, self.channels
        st_in = torch.empty_like(img)
        u8 = torch.empty((h * s, w * s, c), dtype=torch.uint8, device=self.dev) if want_u8 else None
        f32 = torch.empty((c, h * s, w * s), dtype=torch.float32, device=self.dev) if want_f32 else None
        st_in.copy_(img)
        # one eager call first: sizes the workspace (which bumps _ws_gen) and performs the per-device
        # cudaFuncSetAttribute opt-ins, neither of which may happen during capture
        self._restore_call(st_in, fmt, bgr, h, w, t, ov, u8, f32, strict)
        torch.cuda.current_stream(self.dev).synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):  # the capture stream is the current stream inside this block
            self._restore_call(st_in, fmt, bgr, h, w, t, ov, u8, f32, strict)
        g = {"graph": graph, "in": st_in, "u8": u8, "f32": f32, "gen": self._ws_gen}
        while len(self._graphs) >= self.MAX_GRAPHS:  # bounded: every entry pins its static input / output buffers
            self._graphs.pop(next(iter(self._graphs)))
        self._graphs[key] = g
        return g

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
        self._on_device(img)
        img = img.contiguous()
        c = self.channels
        padded = torch.empty((c, plan["hp"], plan["wp"]), dtype=torch.float32, device=self.dev)
        nt = len(plan["ys"]) * len(plan["xs"])
        tiles = torch.empty((nt, plan["th"], plan["tw"], c), dtype=torch.float32, device=self.dev)
        with torch.cuda.device(self.dev):
            check(self.lib.irb_tiles_prepare(self._h, img.data_ptr(), fmt, int(bgr), plan["h"], plan["w"],
                                             0 if tile is None else int(tile), tile_overlap, padded.data_ptr(),
                                             tiles.data_ptr(), self._stream()))
        return tiles

    def tile_batch(self, n: int, th: int, tw: int) -> int:
        """Tiles per forward batch (the "tile_batch" option, else bounded by the 40 GiB workspace cap)."""
        nb = C.c_int()
        check(self.lib.irb_tile_batch(self._h, n, th, tw, C.byref(nb)))
        return nb.value

    def forward_cl(self, tiles: torch.Tensor, out: torch.Tensor = None) -> torch.Tensor:
        """model(x) on channels-last fp32 tiles [n, th, tw, c] -> [n, th*s, tw*s, c], in chunks of tile_batch()
        tiles exactly like irb_restore_image's loop (so a rank's tile range never asks for a larger workspace than
        the single-GPU path). `out` may be a slice of a larger tile buffer."""
        n, th, tw, c = tiles.shape
        s = self.scale
        self._on_device(tiles, out)
        if out is None:
            out = torch.empty((n, th * s, tw * s, c), dtype=torch.float32, device=self.dev)
        assert out.shape == (n, th * s, tw * s, c) and out.is_contiguous() and tiles.is_contiguous()
        if n == 0:
            return out
        need = C.c_size_t()
        with self._lock, torch.cuda.device(self.dev):
            nb = self.tile_batch(n, th, tw)
            check(self.lib.irb_workspace_bytes(self._h, nb, th, tw, C.byref(need)))
            ws = self._workspace(need.value)
            for lo in range(0, n, nb):
                k = min(nb, n - lo)
                check(self.lib.irb_forward_tiles_cl(self._h, tiles[lo:lo + k].data_ptr(), out[lo:lo + k].data_ptr(), k,
                                                    th, tw, ws.data_ptr(), ws.numel(), self._stream()))
        return out

    def stitch(self, tiles_out: torch.Tensor, plan: dict, bgr: bool = False, want_u8=True, want_f32=False):
        from . import ops
        s = plan["scale"]
        with torch.cuda.device(self.dev):
            f32, u8 = ops.stitch(tiles_out, plan["ys"], plan["xs"], s, plan["hp"], plan["wp"], plan["h"] * s,
                                 plan["w"] * s, want_u8=want_u8, reverse_channels=bgr)
        return u8, (f32 if want_f32 else None)

    def stitch_rows(self, tiles_all: torch.Tensor, plan: dict, row0: int, rows: int, bgr: bool = False,
                    want_f32: bool = False, out_u8: torch.Tensor = None):
        """Blend + crop + clamp + round of output rows [row0, row0+rows) (this rank's strip)."""
        from . import ops
        s = plan["scale"]
        with torch.cuda.device(self.dev):
            f32, u8 = ops.stitch_rows(tiles_all, plan["ys"], plan["xs"], s, plan["hp"], plan["wp"], plan["h"] * s,
                                      plan["w"] * s, row0, rows, want_u8=True, want_f32=want_f32,
                                      reverse_channels=bgr, out_u8=out_u8)
        return u8, f32

    def restore(self, img_u8: np.ndarray, tile=None, tile_overlap: int = 32, bgr: bool = False,
                graph: bool = False) -> np.ndarray:
        """Host uint8 [H,W,c] (or [H,W]) in -> host uint8 out; H2D and D2H included."""
        squeeze = img_u8.ndim == 2
        if squeeze:
            img_u8 = img_u8[:, :, None]
        d = torch.from_numpy(np.ascontiguousarray(img_u8)).to(self.dev, non_blocking=True)
        u8, _ = self.restore_device(d, tile, tile_overlap, bgr, graph=graph)
        out = u8.cpu().numpy()
        return out[:, :, 0] if squeeze else out
