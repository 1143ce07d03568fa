#!/usr/bin/env python
"""Drop-in for AgenticIR/executor/*/tools/SwinIR/inference.py (argument list :17-38, output naming :90, :244).

    python inference.py --task {real_sr,color_dn,color_jpeg_car} --model_path P --folder_lq DIR --save_dir DIR
                        [--scale 4 --large_model] [--noise 15] [--jpeg 40] [--tile N --tile_overlap 32]

Every file of --folder_lq is restored by libirb200.so (pad -> tiles -> SwinIR forward -> blend -> crop -> quantise on
the GPU) and written as {save_dir}/output_{suffix}.png with suffix = model_path[model_path.rfind('_')+1:-4], exactly
like get_image_pair() names it. There is no CPU fallback: without a B200 the script exits non-zero.
"""
from __future__ import annotations

import argparse
import glob
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(_HERE)))


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser()
    p.add_argument("--task", type=str, default="color_dn")
    p.add_argument("--scale", type=int, default=1)
    p.add_argument("--noise", type=int, default=15)
    p.add_argument("--jpeg", type=int, default=40)
    p.add_argument("--large_model", action="store_true")
    p.add_argument("--model_path", type=str, default="model_zoo/swinir/001_classicalSR_DIV2K_s48w8_SwinIR-M_x2.pth")
    p.add_argument("--folder_lq", type=str, default=None)
    p.add_argument("--folder_gt", type=str, default=None)
    p.add_argument("--save_dir", type=str, default=None)
    p.add_argument("--tile", type=int, default=None)
    p.add_argument("--tile_overlap", type=int, default=32)
    p.add_argument("--precision", type=str, default=os.environ.get("IRB_PRECISION", "bf16"), choices=["bf16", "fp32"],
                   help="irb200 only: bf16 tensor-core path (default) or the fp32 verification mode")
    return p


def output_name(model_path: str) -> str:
    """inference.py:244: imgname = f"output_{args.model_path[args.model_path.rfind('_')+1:-4]}"."""
    return f"output_{model_path[model_path.rfind('_') + 1:-4]}"


def main(argv=None, state_dict=None, cfg=None) -> int:
    """`state_dict` / `cfg` let tests inject seeded weights (no checkpoints ship with the reference)."""
    args = build_parser().parse_args(argv)
    # inference.py:38
    assert args.task in ["real_sr", "color_dn", "color_jpeg_car"], "task should be real_sr, color_dn or color_jpeg_car"
    import cv2

    from image_restoration_agent_b200 import configs
    from image_restoration_agent_b200.engine import Engine
    from image_restoration_agent_b200.weights import load_checkpoint

    if cfg is None:
        cfg = configs.swinir_config(args.task, scale=args.scale, large_model=args.large_model)
    if state_dict is None:
        keys = ("params_ema",) if args.task == "real_sr" else ("params",)   # define_model() :165, :197-198
        print(f"loading model from {args.model_path}")
        state_dict = load_checkpoint(args.model_path, keys)
    eng = Engine("swinir", cfg, state_dict, precision=args.precision)
    os.makedirs(args.save_dir, exist_ok=True)
    imgname = output_name(args.model_path)
    for path in sorted(glob.glob(os.path.join(args.folder_lq, "*"))):
        img = cv2.imread(path, cv2.IMREAD_COLOR)          # HWC BGR uint8; /255 happens on the device (:240)
        if img is None:
            raise FileNotFoundError(path)
        out = eng.restore(img, args.tile, args.tile_overlap, bgr=True)
        cv2.imwrite(f"{args.save_dir}/{imgname}.png", out)  # :90
    eng.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
