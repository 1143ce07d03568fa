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

import os
import time
from pathlib import Path
from typing import Optional

import numpy as np

from . import configs
from .engine import Engine
from .weights import load_checkpoint

_ENGINES: dict = {}


def get_engine(arch: str, key: str, cfg: dict, state_dict_fn, precision: str = "bf16", device: int = 0) -> Engine:
    k = (arch, key, precision, device)
    if k not in _ENGINES:
        _ENGINES[k] = Engine(arch, cfg, state_dict_fn(), precision=precision, device=device)
    return _ENGINES[k]


def clear_engines():
    for e in _ENGINES.values():
        e.close()
    _ENGINES.clear()


class Tool:
    """Pre/post-conditions of the reference Tool (tool.py:33-62): `input_dir` holds exactly one file, `output_dir`
    is empty before and holds exactly `output.png` after."""

    def __init__(self, tool_name: str, subtask: str):
        self.tool_name = tool_name
        self.subtask = subtask

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
            return load_checkpoint(os.path.join(self.model_zoo, self.model_name), ("params", "params_ema"))
        return get_engine("swinir", self.tool_name + str(id(self.state_dict)), self.cfg, sd, self.precision, self.device)

    def _invoke(self) -> None:
        import cv2

        path = sorted(self.input_dir.glob("*"))[0]
        img = _imread(path, cv2.IMREAD_COLOR)  # uint8 HWC BGR; /255 happens on the device (inference.py:240)
        out = self._engine().restore(img, self.tile, self.tile_overlap, bgr=True)
        cv2.imwrite(str(self.output_dir / "output.png"), out)


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
        return get_engine("restormer", self.opt_task + str(id(self.state_dict)), self.cfg, sd, self.precision,
                          self.device)

    def _invoke(self) -> None:
        import cv2

        path = sorted(self.input_dir.glob("*"))[0]
        gray = self.cfg["inp_channels"] == 1
        img = _imread(path, cv2.IMREAD_GRAYSCALE if gray else cv2.IMREAD_COLOR)  # demo.py:40-47
        # the reference converts BGR->RGB, restores, converts back: same as running on BGR with channel reversal
        out = self._engine().restore(img, self.tile, self.tile_overlap, bgr=not gray)
        cv2.imwrite(str(self.output_dir / "output.png"), out)


def process_image(img_lq: np.ndarray, model: Engine, scale: int, window_size: int, tile=None, tile_overlap=32):
    """Drop-in for model_serving.py:131-175. img_lq: float32 [H, W, 3] BGR in [0,1] (cv2.imread(...)/255) or
    [H, W, 1]; returns uint8 [H*s, W*s, 3] BGR (or [H*s, W*s]). `model` is an irb200 Engine."""
    import torch

    assert model.scale == scale and model.cfg["window_size"] == window_size
    x = np.transpose(img_lq if img_lq.shape[2] == 1 else img_lq[:, :, [2, 1, 0]], (2, 0, 1))
    d = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)).to(f"cuda:{model.device}")
    u8, _ = model.restore_device(d, tile, tile_overlap, bgr=(img_lq.shape[2] == 3))
    out = u8.cpu().numpy()
    return out[:, :, 0] if out.shape[2] == 1 else out


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
        return get_engine("restormer", "svc_" + task + str(id(self.state_dict)), self.cfg, sd, self.precision,
                          self.device)

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
