"""Pins the oracle against the UNMODIFIED reference and writes the golden fixtures under tests/golden/.

Runs only in the build container (needs /root/reference). It
  1. imports the reference nn.Modules (restormer_arch.py via runpy like demo.py:95; network_swinir.py with a
     3-symbol timm.models.layers shim, timm being absent),
  2. loads the seeded synthetic state_dicts of image_restoration_agent_b200.synth into them (load_state_dict),
  3. asserts oracle == reference on every case (fp32, CPU),
  4. stores the reference's inputs/outputs as small .npz fixtures (weights are re-generated from the seed).
The reference tile loop / pad / quantise code lives in scripts that cannot be imported (argparse at module level,
skimage/natsort/flask absent), so for the tool-level fixtures the reference MODULE is driven by oracle/tiling.py.
"""
import os
import runpy
import sys
import tempfile
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = "/root/reference"
RS = f"{REF}/AgenticIR/executor/defocus_deblurring/tools/Restormer"
SW = f"{REF}/AgenticIR/executor/denoising/tools/SwinIR"

from image_restoration_agent_b200 import configs, synth  # noqa: E402
from oracle import restormer_ref, swinir_ref, tiling  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def install_timm_shim():
    timm = types.ModuleType("timm")
    models = types.ModuleType("timm.models")
    layers = types.ModuleType("timm.models.layers")

    class DropPath(torch.nn.Module):  # identity in eval mode
        def __init__(self, drop_prob=0.0):
            super().__init__()

        def forward(self, x):
            return x

    layers.DropPath = DropPath
    layers.to_2tuple = lambda v: tuple(v) if isinstance(v, (tuple, list)) else (v, v)
    layers.trunc_normal_ = lambda t, std=1.0, **kw: torch.nn.init.trunc_normal_(t, std=std)
    sys.modules.update({"timm": timm, "timm.models": models, "timm.models.layers": layers})


def ref_restormer(cfg, sd):
    arch = runpy.run_path(f"{RS}/basicsr/models/archs/restormer_arch.py")
    m = arch["Restormer"](**cfg)
    m.load_state_dict(sd, strict=True)
    return m.eval()


def ref_swinir(cfg, sd):
    install_timm_shim()
    sys.path.insert(0, SW)
    from models.network_swinir import SwinIR  # noqa: E402
    m = SwinIR(**cfg)
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all(k.endswith("relative_position_index") or k.endswith("attn_mask") for k in missing), missing
    return m.eval()


def check(name, a, b, tol=2e-5):
    err = (a - b).abs().max().item()
    print(f"  oracle vs reference [{name}]: max|diff| = {err:.3e}")
    assert err <= tol, (name, err)


def save(name, **arrays):
    path = os.path.join(GOLD, name + ".npz")
    np.savez_compressed(path, **arrays)
    print(f"  wrote {path} ({os.path.getsize(path) / 1024:.1f} KiB)")


def main():
    torch.manual_seed(0)
    torch.set_num_threads(8)
    os.makedirs(GOLD, exist_ok=True)

    # ---------------- Restormer: module level ----------------
    cases = [
        ("restormer_small_withbias", configs.small_restormer_config("WithBias", 3), (1, 3, 32, 40), 1),
        ("restormer_small_biasfree_gray", configs.small_restormer_config("BiasFree", 1), (2, 1, 24, 24), 2),
        ("restormer_full_motion_deblur", configs.restormer_config("Motion_Deblurring"), (1, 3, 64, 64), 0),
        ("restormer_full_real_denoise", configs.restormer_config("Real_Denoising"), (1, 3, 32, 32), 0),
    ]
    for name, cfg, shape, seed in cases:
        print(name)
        sd = synth.restormer_state_dict(cfg, seed=seed)
        ref = ref_restormer(cfg, sd)
        x = torch.rand(shape, generator=torch.Generator().manual_seed(100 + seed))
        with torch.no_grad():
            y_ref = ref(x)
        y_or = restormer_ref.Model(sd, cfg)(x)
        check(name, y_or, y_ref)
        save(name, x=x.numpy(), y=y_ref.numpy(), seed=np.int64(seed))

    # ---------------- Restormer: tool level (pad + tiles + blend + uint8) ----------------
    cfg = configs.small_restormer_config("WithBias", 3)
    sd = synth.restormer_state_dict(cfg, seed=1)
    ref = ref_restormer(cfg, sd)
    img = synth.synthetic_image_u8(45, 61, 3, seed=3)
    for tile, ov in ((None, 32), (24, 8), (32, 16)):
        name = f"restormer_tool_small_t{tile}_o{ov}"
        print(name)
        u8_ref = tiling.restormer_restore_image(img, ref, tile, ov)
        f_ref = tiling.restormer_restore_image(img, ref, tile, ov, return_float=True)
        u8_or = tiling.restormer_restore_image(img, restormer_ref.Model(sd, cfg), tile, ov)
        print(f"  uint8 oracle vs reference: max LSB diff = {np.abs(u8_or.astype(int) - u8_ref.astype(int)).max()}")
        assert np.abs(u8_or.astype(int) - u8_ref.astype(int)).max() <= 1
        save(name, img=img, out_u8=u8_ref, out_f32=f_ref.numpy(), seed=np.int64(1),
             tile=np.int64(-1 if tile is None else tile), overlap=np.int64(ov))

    # ---------------- SwinIR: module level ----------------
    sw_cases = [
        ("swinir_small_color_dn", configs.small_swinir_config("color_dn"), (1, 3, 24, 32), 0),
        ("swinir_small_color_dn_ragged", configs.small_swinir_config("color_dn"), (1, 3, 20, 27), 0),
        ("swinir_small_gray_dn", configs.small_swinir_config("gray_dn"), (2, 1, 16, 16), 1),
        ("swinir_small_jpeg_car_w7", configs.small_swinir_config("color_jpeg_car"), (1, 3, 21, 28), 2),
        ("swinir_small_classical_x4", configs.small_swinir_config("classical_sr", scale=4), (1, 3, 16, 24), 3),
        ("swinir_small_classical_x2", configs.small_swinir_config("classical_sr", scale=2), (1, 3, 16, 16), 3),
        ("swinir_small_realsr_x4", configs.small_swinir_config("real_sr", scale=4), (1, 3, 16, 16), 4),
        ("swinir_small_realsr_large_x4", configs.small_swinir_config("real_sr", scale=4, large_model=True), (1, 3, 16, 16), 5),
        ("swinir_full_gray_dn", configs.swinir_config("gray_dn"), (1, 1, 32, 32), 0),
    ]
    for name, cfg, shape, seed in sw_cases:
        print(name)
        sd = synth.swinir_state_dict(cfg, seed=seed)
        ref = ref_swinir(cfg, sd)
        x = torch.rand(shape, generator=torch.Generator().manual_seed(200 + seed))
        with torch.no_grad():
            y_ref = ref(x)
        y_or = swinir_ref.Model(sd, cfg)(x)
        check(name, y_or, y_ref, tol=5e-5 * max(1.0, cfg["img_range"] / 50))
        save(name, x=x.numpy(), y=y_ref.numpy(), seed=np.int64(seed))

    # ---------------- SwinIR: tool level ----------------
    for task, scale, tile, ov in (("color_dn", 1, None, 32), ("color_dn", 1, 16, 8), ("classical_sr", 4, 16, 8)):
        cfg = configs.small_swinir_config(task, scale=scale)
        sd = synth.swinir_state_dict(cfg, seed=7)
        ref = ref_swinir(cfg, sd)
        img = synth.synthetic_image_u8(30, 27, 3, seed=5)  # "BGR" as cv2.imread would give it
        name = f"swinir_tool_small_{task}_x{scale}_t{tile}_o{ov}"
        print(name)
        u8_ref = tiling.swinir_process_image(img, ref, scale, cfg["window_size"], tile, ov)
        f_ref = tiling.swinir_process_image(img, ref, scale, cfg["window_size"], tile, ov, return_float=True)
        u8_or = tiling.swinir_process_image(img, swinir_ref.Model(sd, cfg), scale, cfg["window_size"], tile, ov)
        d = np.abs(u8_or.astype(int) - u8_ref.astype(int)).max()
        print(f"  uint8 oracle vs reference: max LSB diff = {d}")
        assert d <= 1
        save(name, img=img, out_u8=u8_ref, out_f32=f_ref.numpy(), seed=np.int64(7),
             tile=np.int64(-1 if tile is None else tile), overlap=np.int64(ov), scale=np.int64(scale))

    # ---------------- masks / indices ----------------
    install_timm_shim()
    sys.path.insert(0, SW)
    from models.network_swinir import SwinTransformerBlock, WindowAttention  # noqa: E402
    for ws, (H, W) in ((8, (16, 24)), (7, (14, 21)), (8, (8, 8))):
        blk = SwinTransformerBlock(dim=12, input_resolution=(ws * 8, ws * 8), num_heads=2, window_size=ws, shift_size=ws // 2)
        m_ref = blk.calculate_mask((H, W))
        assert torch.equal(m_ref, swinir_ref.shift_mask(H, W, ws, ws // 2)), "shift mask mismatch"
        wa = WindowAttention(12, (ws, ws), 2)
        assert torch.equal(wa.relative_position_index, swinir_ref.relative_position_index(ws)), "rel index mismatch"
    print("shift masks and relative-position indices: bit-exact")
    print("all golden fixtures written; oracle pinned to the reference modules")


if __name__ == "__main__":
    main()
