"""Golden-fixture case tables shared by the CPU (oracle) and GPU (CUDA path) parity tests."""
from image_restoration_agent_b200 import configs, synth

R_CASES = {
    "restormer_small_withbias": lambda: configs.small_restormer_config("WithBias", 3),
    "restormer_small_biasfree_gray": lambda: configs.small_restormer_config("BiasFree", 1),
    "restormer_full_motion_deblur": lambda: configs.restormer_config("Motion_Deblurring"),
    "restormer_full_real_denoise": lambda: configs.restormer_config("Real_Denoising"),
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
}


def tool_case(name, g):
    """-> (arch, cfg, state_dict, tile, overlap, scale) of a *_tool_* fixture."""
    tile = None if int(g["tile"]) < 0 else int(g["tile"])
    if name.startswith("restormer"):
        cfg = configs.small_restormer_config("WithBias", 3)
        return "restormer", cfg, synth.restormer_state_dict(cfg, seed=int(g["seed"])), tile, int(g["overlap"]), 1
    task = "classical_sr" if "classical_sr" in name else "color_dn"
    scale = int(g["scale"])
    cfg = configs.small_swinir_config(task, scale=scale)
    return "swinir", cfg, synth.swinir_state_dict(cfg, seed=int(g["seed"])), tile, int(g["overlap"]), scale
