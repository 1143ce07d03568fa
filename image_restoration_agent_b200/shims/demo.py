#!/usr/bin/env python
"""Drop-in for AgenticIR/executor/*/tools/Restormer/demo.py (argument list :26-38, file discovery :76-89, output
naming :74, :163-168).

    python demo.py --task {Motion_Deblurring,...,Gaussian_Color_Denoising} --input_dir DIR|FILE --result_dir DIR
                   [--tile N --tile_overlap 32]

Weights are read from the paths demo.py:52-70 uses (relative to the working directory, i.e. the Restormer tool
directory); each image is restored by libirb200.so and written to {result_dir}/{task}/{stem}.png. No CPU fallback.
"""
from __future__ import annotations

import argparse
import os
import re
import sys
from glob import glob

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(_HERE)))

TASKS = ["Motion_Deblurring", "Single_Image_Defocus_Deblurring", "Deraining", "Real_Denoising",
         "Gaussian_Gray_Denoising", "Gaussian_Color_Denoising"]
EXTENSIONS = ["jpg", "JPG", "png", "PNG", "jpeg", "JPEG", "bmp", "BMP"]


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Test Restormer on your own images (irb200 / B200)")
    p.add_argument("--input_dir", default="./demo/degraded/", type=str)
    p.add_argument("--result_dir", default="./demo/restored/", type=str)
    p.add_argument("--task", required=True, type=str, choices=TASKS)
    p.add_argument("--tile", type=int, default=None)
    p.add_argument("--tile_overlap", type=int, default=32)
    p.add_argument("--precision", type=str, default=os.environ.get("IRB_PRECISION", "bf16"), choices=["bf16", "fp32"])
    return p


def natural_key(s: str):
    """natsorted() ordering of demo.py:86 (natsort is not a dependency here): digit runs compare as integers."""
    return [int(t) if t.isdigit() else t for t in re.split(r"(\d+)", s)]


def find_files(inp_dir: str):
    if any(inp_dir.endswith(ext) for ext in EXTENSIONS):
        return [inp_dir]
    files = []
    for ext in EXTENSIONS:
        files.extend(glob(os.path.join(inp_dir, "*." + ext)))
    return sorted(files, key=natural_key)


def main(argv=None, state_dict=None, cfg=None) -> int:
    args = build_parser().parse_args(argv)
    import cv2

    from image_restoration_agent_b200 import configs
    from image_restoration_agent_b200.engine import Engine
    from image_restoration_agent_b200.tool_api import RestormerB200
    from image_restoration_agent_b200.weights import load_checkpoint

    task = args.task
    out_dir = os.path.join(args.result_dir, task)
    os.makedirs(out_dir, exist_ok=True)
    files = find_files(args.input_dir)
    if len(files) == 0:
        raise Exception(f"No files found at {args.input_dir}")
    if cfg is None:
        cfg = configs.restormer_config(task)
    if state_dict is None:
        weights = RestormerB200.WEIGHTS[task]
        print(f"\n ==> Running {task} with weights {weights}\n ")
        state_dict = load_checkpoint(weights, ("params",))
    eng = Engine("restormer", cfg, state_dict, precision=args.precision)
    gray = task == "Gaussian_Gray_Denoising"
    for file_ in files:
        img = cv2.imread(file_, cv2.IMREAD_GRAYSCALE if gray else cv2.IMREAD_COLOR)
        if img is None:
            raise FileNotFoundError(file_)
        # load_img converts BGR->RGB and save_img converts back (:40-44): restore on BGR with channel reversal
        out = eng.restore(img, args.tile, args.tile_overlap, bgr=not gray)
        stem = os.path.splitext(os.path.split(file_)[-1])[0]
        cv2.imwrite(os.path.join(out_dir, stem + ".png"), out)
    print(f"\nRestored images are saved at {out_dir}")
    eng.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
