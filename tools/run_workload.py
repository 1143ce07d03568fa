"""Runs one BASELINE workload N times through the Engine (for ncu launch lists / quick timing).

usage: python tools/run_workload.py {cfg2|cfg5|swinir_dn|swinir_sr} [iters] [tile_batch]
"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from image_restoration_agent_b200 import configs, synth  # noqa: E402
from image_restoration_agent_b200.engine import Engine  # noqa: E402

WORKLOADS = {
    # name: (arch, cfg, (h, w, c), tile, overlap)
    "cfg2": ("restormer", lambda: configs.restormer_config("Motion_Deblurring"), (720, 1280, 3), 256, 32),
    "cfg5": ("restormer", lambda: configs.restormer_config("Real_Denoising"), (2160, 3840, 3), 512, 64),
    "swinir_dn": ("swinir", lambda: configs.swinir_config("gray_dn"), (256, 256, 1), None, 32),
    "swinir_sr": ("swinir", lambda: configs.swinir_config("classical_sr", scale=4), (512, 512, 3), 256, 32),
}


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    tb = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    arch, mk, (h, w, c), tile, ov = WORKLOADS[name]
    cfg = mk()
    sd = (synth.restormer_state_dict if arch == "restormer" else synth.swinir_state_dict)(cfg, seed=0)
    eng = Engine(arch, cfg, sd, precision="bf16")
    eng.set_tile_batch(tb)
    img = torch.from_numpy(synth.synthetic_image_u8(h, w, c, seed=0)).cuda()
    torch.cuda.synchronize()
    for i in range(iters):
        t0 = time.time()
        eng.restore_device(img, tile, ov)
        torch.cuda.synchronize()
        print(f"{name} iter {i}: {(time.time() - t0) * 1e3:.2f} ms", flush=True)


if __name__ == "__main__":
    main()
