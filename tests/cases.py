"""Golden-fixture case tables shared by the CPU (oracle) and GPU (CUDA path) parity tests."""
from image_restoration_agent_b200 import configs, synth

R_CASES = {
    "restormer_small_withbias": lambda: configs.small_restormer_config("WithBias", 3),
    "restormer_small_biasfree_gray": lambda: configs.small_restormer_config("BiasFree", 1),
    "restormer_full_motion_deblur": lambda: configs.restormer_config("Motion_Deblurring"),
    "restormer_full_real_denoise": lambda: configs.restormer_config("Real_Denoising"),
    # round 2: one tile of the benchmarked configurations (MDTA Gram K = 65 536 / 262 144), tools/make_golden_full.py
    "restormer_cfg2_tile256": lambda: configs.restormer_config("Motion_Deblurring"),
    "restormer_cfg5_tile512": lambda: configs.restormer_config("Real_Denoising"),
}
X_CASES = {  # X-Restormer, tools/make_golden_full.py xrestormer
    "xrestormer_small_withbias": lambda: configs.small_xrestormer_config("WithBias"),
    "xrestormer_small_biasfree": lambda: configs.small_xrestormer_config("BiasFree"),
    "xrestormer_full": lambda: configs.xrestormer_config(),
    "xrestormer_small_sr_x4": lambda: configs.small_xrestormer_config("WithBias", scale=4),
}
H_CASES = {  # HAT, tools/make_golden_full.py hat
    "hat_small": lambda: configs.small_hat_config(),
    "hat_small_w8_x2": lambda: configs.small_hat_config(window_size=8, upscale=2),
    "hat_full": lambda: configs.hat_config(),
}
S_CASES = {
    "swinir_small_color_dn": lambda: configs.small_swinir_config("color_dn"),
    "swinir_small_color_dn_ragged": lambda: configs.small_swinir_config("color_dn"),
    "swinir_small_gray_dn": lambda: configs.small_swinir_config("gray_dn"),
    "swinir_small_jpeg_car_w7": lambda: configs.small_swinir_config("color_jpeg_car"),
    "swinir_small_classical_x4": lambda: configs.small_swinir_config("classical_sr", scale=4),
    "swinir_small_classical_x2": lambda: configs.small_swinir_config("classical_sr", scale=2),
    "swinir_small_realsr_x4": lambda: configs.small_swinir_config("real_sr", scale=4),
    "swinir_small_realsr_large_x4": lambda: configs.small_swinir_config("real_sr", scale=4, large_model=True),
    "swinir_full_gray_dn": lambda: configs.swinir_config("gray_dn"),
    # round 2: full depth (36 / 54 STL) colour and SR models on 64x64 image-like inputs
    "swinir_full_color_dn": lambda: configs.swinir_config("color_dn"),
    "swinir_full_classical_x4": lambda: configs.swinir_config("classical_sr", scale=4),
    "swinir_full_realsr_large_x4": lambda: configs.swinir_config("real_sr", scale=4, large_model=True),
    # lightweight SR: the 'pixelshuffledirect' upsampler (x2, x3 with a ragged 21x30 input, full depth x4)
    "swinir_small_lightweight_x2": lambda: configs.small_swinir_config("lightweight_sr", scale=2),
    "swinir_small_lightweight_x3": lambda: configs.small_swinir_config("lightweight_sr", scale=3),
    "swinir_full_lightweight_x4": lambda: configs.swinir_config("lightweight_sr", scale=4),
    # DiffBIR's stage-1 SwinIR: PixelUnshuffle(8) head, nearest+conv x8 tail; from the imported
    # DiffBIR/model/swinir.py (tools/make_golden_full.py diffbir)
    "diffbir_swinir_small": lambda: configs.small_diffbir_swinir_config(),
    "diffbir_swinir_full": lambda: configs.diffbir_swinir_config(),
}


def tool_case(name, g):
    """-> (arch, cfg, state_dict, tile, overlap, scale) of a *_tool_* fixture."""
    tile = None if int(g["tile"]) < 0 else int(g["tile"])
    if name.startswith("restormer"):
        cfg = (configs.restormer_config("Motion_Deblurring") if "cfg2crop" in name
               else configs.small_restormer_config("WithBias", 3))
        return "restormer", cfg, synth.restormer_state_dict(cfg, seed=int(g["seed"])), tile, int(g["overlap"]), 1
    task = "classical_sr" if "classical_sr" in name else "color_dn"
    scale = int(g["scale"])
    full = "full" in g.files and int(g["full"]) == 1
    cfg = configs.swinir_config(task, scale=scale) if full else configs.small_swinir_config(task, scale=scale)
    return "swinir", cfg, synth.swinir_state_dict(cfg, seed=int(g["seed"])), tile, int(g["overlap"]), scale
