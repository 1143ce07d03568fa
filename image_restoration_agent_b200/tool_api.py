"""Host-side mirror of the reference's tool / service interface for the SwinIR and Restormer tools.

Same names, argument meaning and error behaviour as the reference, with the tiled forward pass served by
libirb200.so instead of a `conda run ... python inference.py` subprocess:

  Tool.__call__(input_dir, output_dir, silent)          AgenticIR/executor/tool.py:33-62
  SwinIR(subtask, pretrained_on) / Restormer(subtask)   AgenticIR/executor/multitask_tools.py:81-173
  process_image(img_lq, model, scale, window_size, ...) AgentApp/model_service/denoising/SwinIR/model_serving.py:131-175
  RestormerService.restore_image(img_bytes, task, ...)  AgentApp/model_service/motion_deblurring/Restormer/model_serving.py:157-208

Models stay resident across calls: engines are cached per (tool, precision, device).
INTEGRATION.md shows the three-line change that registers these classes in the reference's toolboxes.
"""
from __future__ import annotations

import hashlib
import json
import os
import threading
import time
from pathlib import Path
from typing import Optional

import numpy as np

from . import configs
from .engine import Engine
from .weights import load_checkpoint

_ENGINES: dict = {}
_ENGINES_LOCK = threading.Lock()


def _cfg_digest(cfg: dict) -> str:
    return hashlib.sha1(json.dumps(cfg, sort_keys=True, default=str).encode()).hexdigest()[:16]


def get_engine(arch: str, source, cfg: dict, state_dict_fn, precision: str = "bf16", device: int = 0) -> Engine:
    """Resident engines, one per (arch, weights source, config, precision, device). `source` is the resolved
    checkpoint path (str) or the caller's in-memory state_dict object; for the latter the cache entry keeps a
    reference to it, so its id() cannot be recycled for other weights while the entry lives. Thread-safe: Flask
    worker threads racing on a cold cache build one engine."""
    src_key = ("path", os.path.abspath(source)) if isinstance(source, (str, os.PathLike)) else ("obj", id(source))
    k = (arch, src_key, _cfg_digest(cfg), precision, int(device))
    with _ENGINES_LOCK:
        hit = _ENGINES.get(k)
        if hit is None:
            hit = (Engine(arch, cfg, state_dict_fn(), precision=precision, device=device), source)
            _ENGINES[k] = hit
        return hit[0]


def clear_engines():
    with _ENGINES_LOCK:
        for e, _ in _ENGINES.values():
            e.close()
        _ENGINES.clear()


class _LocalTool:
    """Mirror of the reference Tool's pre/post-conditions (tool.py:33-62), used only when the reference package is
    not importable: `input_dir` holds exactly one file, `output_dir` is empty before and holds exactly `output.png`
    after."""

    def __init__(self, tool_name: str, subtask: str, work_dir=None, script_rel_path=None):
        self.tool_name = tool_name
        self.subtask = subtask
        self.work_dir = None
        self.script_path = None

    def __call__(self, input_dir: Path, output_dir: Path, silent: bool = False, *args) -> None:
        input_dir, output_dir = Path(input_dir), Path(output_dir)
        if not silent:
            print("-" * 100)
            print(f"Subtask\t: {self.subtask}")
            print(f"Tool\t: {self.tool_name}")
            print(f"Input\t: {list(input_dir.glob('*'))[0]}")
        start = time.time()
        self.input_dir, self.output_dir = input_dir, output_dir
        self._precheck()
        self._invoke(*args)
        self._postcheck()
        if not silent:
            print(f"Output\t: {list(output_dir.glob('*'))[0]}")
            print(f"Time\t: {round(time.time() - start, 3)}s")

    def _precheck(self) -> None:
        assert len(os.listdir(self.input_dir)) == 1, "The input directory should contain the input only."
        assert os.listdir(self.output_dir) == [], "The output directory should be empty."

    def _postcheck(self) -> None:
        output = list(self.output_dir.glob("*"))
        assert len(output) == 1, "There're other files in the same directory as the output image."
        if output[0].name != "output.png":
            output[0].replace(self.output_dir / "output.png")

    def _invoke(self, *args) -> None:
        raise NotImplementedError


def _reference_tool_base():
    """Inside the reference process (AgenticIR/ on sys.path) the B200 tools subclass the reference's own
    executor.tool.Tool, so isinstance checks and the toolbox lists (`list[Tool]`) accept them unchanged."""
    try:
        from executor.tool import Tool as RefTool  # noqa: PLC0415
        return RefTool
    except Exception:  # noqa: BLE001  (package absent, or its heavy imports unavailable)
        return None


Tool = _reference_tool_base() or _LocalTool


def _imread(path, flags):
    import cv2

    img = cv2.imread(str(path), flags)
    if img is None:
        raise FileNotFoundError(f"cannot read image {path}")
    return img


class SwinIRB200(Tool):
    """In-process replacement of multitask_tools.SwinIR (same tool_name: swinir_15 / _50 / _gan / _psnr / _40)."""

    OPT = {  # multitask_tools.py:98-118
        "super_resolution": ("real_sr", {"gan": "003_realSR_BSRGAN_DFOWMFC_s64w8_SwinIR-L_x4_GAN.pth",
                                         "psnr": "003_realSR_BSRGAN_DFOWMFC_s64w8_SwinIR-L_x4_PSNR.pth"}),
        "denoising": ("color_dn", {"15": "005_colorDN_DFWB_s128w8_SwinIR-M_noise15.pth",
                                   "50": "005_colorDN_DFWB_s128w8_SwinIR-M_noise50.pth"}),
        "jpeg_compression_artifact_removal": ("color_jpeg_car", {"40": "006_colorCAR_DFWB_s126w7_SwinIR-M_jpeg40.pth"}),
    }

    def __init__(self, subtask: str, pretrained_on: str, model_zoo: Optional[str] = None, state_dict=None,
                 precision: str = "bf16", device: int = 0, tile: Optional[int] = None, tile_overlap: int = 32,
                 cfg: Optional[dict] = None):
        super().__init__(tool_name=f"swinir_{pretrained_on}", subtask=subtask)
        self.opt_task, names = self.OPT[subtask]
        self.model_name = names[pretrained_on]
        self.scale = 4 if subtask == "super_resolution" else 1  # multitask_tools.py:128-132: --scale 4 --large_model
        self.cfg = cfg or configs.swinir_config(self.opt_task, scale=self.scale,
                                                large_model=(subtask == "super_resolution"))
        self.model_zoo, self.state_dict = model_zoo, state_dict
        self.precision, self.device = precision, device
        self.tile, self.tile_overlap = tile, tile_overlap  # the executor never passes --tile

    def _engine(self) -> Engine:
        def sd():
            if self.state_dict is not None:
                return self.state_dict
            # define_model(): 'params_ema' for real_sr, 'params' otherwise, bare dict if absent (inference.py:165,197-198)
            keys = ("params_ema",) if self.opt_task == "real_sr" else ("params",)
            return load_checkpoint(os.path.join(self.model_zoo, self.model_name), keys)
        src = self.state_dict if self.state_dict is not None else os.path.join(self.model_zoo, self.model_name)
        return get_engine("swinir", src, self.cfg, sd, self.precision, self.device)

    def restore_u8_device(self, img_bgr, graph: bool = False):
        """uint8 [H,W,3] BGR on the GPU (what cv2.imread of input.png holds) -> uint8 [sH,sW,3] BGR on the GPU (what
        output.png would hold): the tool body without the file system (Chain keeps the hand-off in HBM)."""
        return self._engine().restore_device(img_bgr, self.tile, self.tile_overlap, bgr=True, graph=graph)[0]

    def _invoke(self) -> None:
        import cv2

        path = sorted(Path(self.input_dir).glob("*"))[0]
        img = _imread(path, cv2.IMREAD_COLOR)  # uint8 HWC BGR; /255 happens on the device (inference.py:240)
        out = self._engine().restore(img, self.tile, self.tile_overlap, bgr=True)
        cv2.imwrite(str(Path(self.output_dir) / "output.png"), out)


class RestormerB200(Tool):
    """In-process replacement of multitask_tools.Restormer (tool_name 'restormer')."""

    OPT = {"denoising": "Real_Denoising", "motion_deblurring": "Motion_Deblurring",
           "defocus_deblurring": "Single_Image_Defocus_Deblurring", "deraining": "Deraining"}  # :150-155
    WEIGHTS = {  # demo.py:52-70
        "Motion_Deblurring": "Motion_Deblurring/pretrained_models/motion_deblurring.pth",
        "Single_Image_Defocus_Deblurring": "Defocus_Deblurring/pretrained_models/single_image_defocus_deblurring.pth",
        "Deraining": "Deraining/pretrained_models/deraining.pth",
        "Real_Denoising": "Denoising/pretrained_models/real_denoising.pth",
        "Gaussian_Color_Denoising": "Denoising/pretrained_models/gaussian_color_denoising_blind.pth",
        "Gaussian_Gray_Denoising": "Denoising/pretrained_models/gaussian_gray_denoising_blind.pth",
    }

    def __init__(self, subtask: str, weights_root: Optional[str] = None, state_dict=None, precision: str = "bf16",
                 device: int = 0, tile: Optional[int] = None, tile_overlap: int = 32, cfg: Optional[dict] = None):
        super().__init__(tool_name="restormer", subtask=subtask)
        self.opt_task = self.OPT[subtask]
        self.cfg = cfg or configs.restormer_config(self.opt_task)
        self.weights_root, self.state_dict = weights_root, state_dict
        self.precision, self.device = precision, device
        self.tile, self.tile_overlap = tile, tile_overlap

    def _engine(self) -> Engine:
        def sd():
            if self.state_dict is not None:
                return self.state_dict
            return load_checkpoint(os.path.join(self.weights_root, self.WEIGHTS[self.opt_task]), ("params",))
        src = (self.state_dict if self.state_dict is not None
               else os.path.join(self.weights_root, self.WEIGHTS[self.opt_task]))
        return get_engine("restormer", src, self.cfg, sd, self.precision, self.device)

    def restore_u8_device(self, img_bgr, graph: bool = False):
        """uint8 [H,W,3] BGR (or [H,W,1] gray) on the GPU -> restored uint8 of the same layout on the GPU."""
        gray = self.cfg["inp_channels"] == 1
        return self._engine().restore_device(img_bgr, self.tile, self.tile_overlap, bgr=not gray, graph=graph)[0]

    def _invoke(self) -> None:
        import cv2

        path = sorted(Path(self.input_dir).glob("*"))[0]
        gray = self.cfg["inp_channels"] == 1
        img = _imread(path, cv2.IMREAD_GRAYSCALE if gray else cv2.IMREAD_COLOR)  # demo.py:40-47
        # the reference converts BGR->RGB, restores, converts back: same as running on BGR with channel reversal
        out = self._engine().restore(img, self.tile, self.tile_overlap, bgr=not gray)
        cv2.imwrite(str(Path(self.output_dir) / "output.png"), out)


class XRestormerB200(Tool):
    """In-process replacement of multitask_tools.XRestormer (tool_name 'xrestormer', a BasicSRModel tool:
    multitask_tools.py:67-78) for its five subtasks. The reference runs BasicSR's test pipeline on the yml of
    executor/<subtask>/configs/xrestormer.yml: reflect pad to a multiple of 64 (16 for the scale-4 SR config, whose
    network first up-samples the input bilinearly), whole image, crop, tensor2img."""

    SUBTASKS = configs.XRESTORMER_SUBTASKS
    WEIGHTS = {"dehazing": "dehaze_300k.pth", "denoising": "denoise_300k.pth", "deraining": "derain_155k.pth",
               "motion_deblurring": "deblur_300k.pth", "super_resolution": "sr_300k.pth"}  # path.pretrain_network_g

    def __init__(self, subtask: str, weights_root: Optional[str] = None, state_dict=None, precision: str = "bf16",
                 device: int = 0, cfg: Optional[dict] = None):
        if subtask not in self.SUBTASKS:
            raise ValueError(f"irb200 X-Restormer serves {self.SUBTASKS}; got {subtask!r}")
        super().__init__(tool_name="xrestormer", subtask=subtask)
        self.cfg = cfg or configs.xrestormer_config(subtask)
        self.weights_root, self.state_dict = weights_root, state_dict
        self.precision, self.device = precision, device

    def _engine(self) -> Engine:
        def sd():
            if self.state_dict is not None:
                return self.state_dict
            return load_checkpoint(os.path.join(self.weights_root, self.WEIGHTS[self.subtask]), ("params_ema", "params"))
        src = (self.state_dict if self.state_dict is not None
               else os.path.join(self.weights_root, self.WEIGHTS[self.subtask]))
        return get_engine("xrestormer", src, self.cfg, sd, self.precision, self.device)

    def restore_u8_device(self, img_bgr, graph: bool = False):
        return self._engine().restore_device(img_bgr, None, 0, bgr=True, graph=graph)[0]

    def _invoke(self) -> None:
        import cv2

        path = sorted(Path(self.input_dir).glob("*"))[0]
        img = _imread(path, cv2.IMREAD_COLOR)
        out = self._engine().restore(img, None, 0, bgr=True)
        cv2.imwrite(str(Path(self.output_dir) / "output.png"), out)


class HATB200(Tool):
    """In-process replacement of the agent's HAT tool (executor/super_resolution/__init__.py:12-20: tool_name 'hat', a
    BasicSRModel tool running hat/test.py on executor/super_resolution/configs/hat.yml). BasicSR's test pipeline with
    model_type HATModel (hat/models/hat_model.py:16-109): cv2.imread BGR -> RGB / 255, reflect pad to the window (16),
    tile_size 512 / tile_pad 32 crop-with-pad tiles pasted without blending, crop, tensor2img -> output.png.
    Checkpoint key 'params_ema' (hat.yml path.param_key_g)."""

    WEIGHTS = "Real_HAT_GAN_SRx4.pth"  # hat.yml path.pretrain_network_g

    def __init__(self, weights_root: Optional[str] = None, state_dict=None, precision: str = "bf16", device: int = 0,
                 cfg: Optional[dict] = None, tile: Optional[dict] = None):
        super().__init__(tool_name="hat", subtask="super_resolution")
        self.cfg = cfg or configs.hat_config()
        self.tile = dict(configs.HAT_TILE if tile is None else tile)   # {} = HATModel.process (no 'tile' key in the yml)
        self.weights_root, self.state_dict = weights_root, state_dict
        self.precision, self.device = precision, device

    def _engine(self) -> Engine:
        def sd():
            if self.state_dict is not None:
                return self.state_dict
            return load_checkpoint(os.path.join(self.weights_root, self.WEIGHTS), ("params_ema", "params"))
        src = self.state_dict if self.state_dict is not None else os.path.join(self.weights_root, self.WEIGHTS)
        return get_engine("hat", src, self.cfg, sd, self.precision, self.device)

    def restore_u8_device(self, img_bgr, graph: bool = False):
        return self._engine().restore_crop_device(img_bgr, self.tile.get("tile_size"), self.tile.get("tile_pad", 0),
                                                  bgr=True)[0]

    def _invoke(self) -> None:
        import cv2

        path = sorted(Path(self.input_dir).glob("*"))[0]
        img = _imread(path, cv2.IMREAD_COLOR)
        out = self._engine().restore_crop(img, self.tile.get("tile_size"), self.tile.get("tile_pad", 0), bgr=True)
        cv2.imwrite(str(Path(self.output_dir) / "output.png"), out)


class Chain:
    """A plan of several tools run back to back on one image (BASELINE config 4: derain -> denoise -> x4 SR) with the
    inter-stage hand-off kept in HBM.

    The reference chains tools through files: every tool writes a uint8 `output.png`, which is copied to the next
    tool's `input.png` (AgenticIR/executor/tool.py:56-62, AgentApp/agentic_api.py:200-210), so the image is QUANTISED
    to uint8 between stages. PNG is lossless, hence the next stage sees exactly the uint8 array the
    previous one produced: here that array simply stays on the GPU. Decode once, H2D once, D2H once, encode once;
    the bytes equal the file-based chain's."""

    def __init__(self, tools):
        self.tools = list(tools)
        assert self.tools, "empty chain"

    def run_device(self, img_bgr_u8, graph: bool = False):
        x = img_bgr_u8
        for k, tool in enumerate(self.tools):
            x = tool.restore_u8_device(x, graph=graph)
            if graph and k + 1 < len(self.tools):
                # a graph's output buffer is static: the next call of the SAME tool would overwrite it in place
                # only when one tool appears twice in a chain
                if any(t is tool for t in self.tools[k + 1:]):
                    x = x.clone()
        return x

    def run(self, img_bgr_u8: np.ndarray, graph: bool = False) -> np.ndarray:
        import torch

        dev = f"cuda:{self.tools[0].device}"
        d = torch.from_numpy(np.ascontiguousarray(img_bgr_u8)).to(dev, non_blocking=True)
        return self.run_device(d, graph=graph).cpu().numpy()

    def __call__(self, input_dir: Path, output_dir: Path, silent: bool = False) -> None:
        """Same contract as Tool.__call__ (one file in, `output.png` out) for the whole chain."""
        import cv2

        input_dir, output_dir = Path(input_dir), Path(output_dir)
        assert len(os.listdir(input_dir)) == 1, "The input directory should contain the input only."
        assert os.listdir(output_dir) == [], "The output directory should be empty."
        start = time.time()
        img = _imread(sorted(input_dir.glob("*"))[0], cv2.IMREAD_COLOR)
        out = self.run(img)
        cv2.imwrite(str(output_dir / "output.png"), out)
        if not silent:
            print(f"Chain\t: {' -> '.join(t.tool_name for t in self.tools)}")
            print(f"Output\t: {output_dir / 'output.png'}")
            print(f"Time\t: {round(time.time() - start, 3)}s")


def process_image(img_lq: np.ndarray, model: Engine, scale: int, window_size: int, tile=None, tile_overlap=32):
    """Drop-in for model_serving.py:131-175. img_lq: float32 [H, W, 3] BGR in [0,1] (cv2.imread(...)/255) or
    [H, W, 1]; returns uint8 [H*s, W*s, 3] BGR (or [H*s, W*s]). `model` is an irb200 Engine."""
    import torch

    assert model.scale == scale and model.cfg["window_size"] == window_size
    x = np.transpose(img_lq if img_lq.shape[2] == 1 else img_lq[:, :, [2, 1, 0]], (2, 0, 1))
    d = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).to(f"cuda:{model.device}")
    # model_serving.py:146-166 has no `tile % window_size` assert (the option is applied under the engine's lock)
    u8, _ = model.restore_device(d, tile, tile_overlap, bgr=(img_lq.shape[2] == 3), strict_tile=False)
    out = u8.cpu().numpy()
    return out[:, :, 0] if out.shape[2] == 1 else out


def process_images(imgs_lq, model: Engine, scale: int, window_size: int, tile=None, tile_overlap=32):
    """The `/predict_batch` loop of the SwinIR service (model_serving.py:285-306: process_image per uploaded file) on a
    list of float32 BGR images; the engine stays resident and same-shaped images reuse one workspace (and one CUDA
    graph when the caller passes graph-enabled engines). Under an initialised torch.distributed group the list is
    sharded over the ranks (parallel.partition) and rank 0 receives every result."""
    import torch.distributed as dist

    from . import parallel
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return [process_image(im, model, scale, window_size, tile, tile_overlap) for im in imgs_lq]
    rank, world = dist.get_rank(), dist.get_world_size()
    lo, hi = parallel.partition(len(imgs_lq), world)[rank]
    mine = [process_image(imgs_lq[i], model, scale, window_size, tile, tile_overlap) for i in range(lo, hi)]
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(mine, gathered, dst=0)
    return [o for part in gathered for o in part] if rank == 0 else None


class RestormerService:
    """Drop-in for the Flask wrapper's service object (model_serving.py:55-208): load_model + restore_image."""

    def __init__(self, task: str, weights_dir: Optional[str] = None, state_dict=None, precision: str = "bf16",
                 device: int = 0, cfg: Optional[dict] = None):
        self.task, self.weights_dir, self.state_dict = task, weights_dir, state_dict
        self.precision, self.device = precision, device
        self.cfg = cfg or configs.restormer_config(task)
        self.img_multiple_of = 8
        self.model = self.load_model(task)

    def load_model(self, task: str) -> Engine:
        if task != self.task:
            raise ValueError(f"Not supported task {task}")

        def sd():
            if self.state_dict is not None:
                return self.state_dict
            name = os.path.basename(RestormerB200.WEIGHTS[task])
            return load_checkpoint(os.path.join(self.weights_dir, name), ("params",))
        src = (self.state_dict if self.state_dict is not None
               else os.path.join(self.weights_dir, os.path.basename(RestormerB200.WEIGHTS[task])))
        return get_engine("restormer", src, self.cfg, sd, self.precision, self.device)

    def restore_image(self, img_bytes: bytes, task: str, tile=None, tile_overlap=32) -> bytes:
        import cv2

        is_gray = task == "Gaussian_Gray_Denoising"
        arr = np.frombuffer(img_bytes, np.uint8)
        img = cv2.imdecode(arr, cv2.IMREAD_GRAYSCALE if is_gray else cv2.IMREAD_COLOR)
        if img is None:
            raise ValueError("cannot decode image")
        out = self.model.restore(img, tile, tile_overlap, bgr=not is_gray)
        ok, enc = cv2.imencode(".png", out)
        if not ok:
            raise ValueError("Img_to_bytes failed")
        return enc.tobytes()
