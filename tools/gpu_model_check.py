"""Model-level bring-up on a B200: Engine vs the committed golden fixtures (reference outputs), fp32 and bf16,
plus a first timing of the BASELINE config-2 workload. Development tool; the judged tests live in tests/.
"""
import glob
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from image_restoration_agent_b200 import configs, synth  # noqa: E402
from image_restoration_agent_b200.engine import Engine  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
RES = {}

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


def psnr(a, b):
    mse = float(((a - b) ** 2).mean())
    return 99.0 if mse == 0 else 10 * np.log10(1.0 / mse)


def module_cases(arch, cases):
    for name, mk in cases.items():
        g = np.load(os.path.join(GOLD, name + ".npz"))
        cfg = mk()
        sd = (synth.restormer_state_dict if arch == "restormer" else synth.swinir_state_dict)(cfg, seed=int(g["seed"]))
        x = torch.from_numpy(g["x"]).cuda()
        for prec, tol in (("fp32", 1e-4), ("bf16", 2e-2)):
            try:
                eng = Engine(arch, cfg, sd, precision=prec)
                y = eng(x).cpu().numpy()
                err = float(np.abs(y - g["y"]).max())
                p = psnr(y, g["y"])
                ok = err <= tol and (prec == "fp32" or p >= 45.0)
                RES[f"{name}[{prec}]"] = {"max_abs": err, "psnr": p, "ok": bool(ok)}
                print(f"{'OK ' if ok else 'BAD'} {name}[{prec}] max_abs={err:.3e} psnr={p:.1f} dB", flush=True)
                eng.close()
            except Exception as e:  # noqa: BLE001
                RES[f"{name}[{prec}]"] = {"ok": False, "exception": repr(e)}
                print(f"EXC {name}[{prec}]: {e!r}", flush=True)
                torch.cuda.synchronize()


def tool_cases(prefix=""):
    for path in sorted(glob.glob(os.path.join(GOLD, prefix + "*_tool_*.npz"))):
        name = os.path.basename(path)[:-4]
        g = np.load(path)
        arch = "restormer" if name.startswith("restormer") else "swinir"
        if arch == "restormer":
            cfg = configs.small_restormer_config("WithBias", 3)
            sd = synth.restormer_state_dict(cfg, seed=int(g["seed"]))
        else:
            task = "classical_sr" if "classical_sr" in name else "color_dn"
            cfg = configs.small_swinir_config(task, scale=int(g["scale"]))
            sd = synth.swinir_state_dict(cfg, seed=int(g["seed"]))
        tile = None if int(g["tile"]) < 0 else int(g["tile"])
        for prec, tol in (("fp32", 1e-4), ("bf16", 2e-2)):
            try:
                eng = Engine(arch, cfg, sd, precision=prec)
                d = torch.from_numpy(g["img"]).cuda()
                u8, f32 = eng.restore_device(d, tile, int(g["overlap"]), bgr=(arch == "swinir"), want_f32=True)
                f = f32.cpu().numpy()
                ref_f = g["out_f32"]
                if arch == "swinir":
                    pass
                err = float(np.abs(f - ref_f).max())
                lsb = int(np.abs(u8.cpu().numpy().astype(int) - g["out_u8"].astype(int)).max())
                ok = err <= tol and lsb <= 1
                RES[f"{name}[{prec}]"] = {"max_abs": err, "lsb": lsb, "ok": bool(ok)}
                print(f"{'OK ' if ok else 'BAD'} {name}[{prec}] max_abs={err:.3e} u8 max LSB diff={lsb}", flush=True)
                eng.close()
            except Exception as e:  # noqa: BLE001
                RES[f"{name}[{prec}]"] = {"ok": False, "exception": repr(e)}
                print(f"EXC {name}[{prec}]: {e!r}", flush=True)
                torch.cuda.synchronize()


def time_cfg2():
    cfg = configs.restormer_config("Motion_Deblurring")
    sd = synth.restormer_state_dict(cfg, seed=0)
    eng = Engine("restormer", cfg, sd, precision="bf16")
    img = torch.from_numpy(synth.synthetic_image_u8(720, 1280, 3, seed=0)).cuda()
    for tb in (0, 8, 4):
        eng.set_tile_batch(tb)
        for _ in range(2):
            eng.restore_device(img, 256, 32)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 3
        for _ in range(n):
            eng.restore_device(img, 256, 32)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        RES[f"cfg2_time_tile_batch{tb}"] = {"ms": ms, "mp_per_s": 0.9216 / (ms / 1e3), "ok": True}
        print(f"cfg2 1280x720 t256/o32 bf16 tile_batch={tb}: {ms:.2f} ms -> {0.9216 / (ms / 1e3):.2f} MP/s", flush=True)
    eng.close()


if __name__ == "__main__":
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    which = sys.argv[1:] or ["restormer", "swinir", "tools", "time"]
    t0 = time.time()
    if "restormer" in which:
        module_cases("restormer", R_CASES)
    if "swinir" in which:
        module_cases("swinir", S_CASES)
    if "tools" in which:
        tool_cases()
    if "tools_restormer" in which:
        tool_cases("restormer")
    if "time" in which:
        time_cfg2()
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/model_check_" + "+".join(which) + ".json", "w") as f:
        json.dump(RES, f, indent=1)
    nbad = sum(1 for v in RES.values() if not v.get("ok"))
    print(f"done in {time.time() - t0:.1f}s: {len(RES)} checks, {nbad} bad")
    sys.exit(1 if nbad else 0)
