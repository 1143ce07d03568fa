#!/usr/bin/env python
"""bench.py — restored megapixels/s of the tiled restoration tool body on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg5|cfg1|cfg3|cfg4|...] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one batch of the workload: pad -> tiles -> model forward on every tile
-> overlap blend -> crop -> clamp -> round -> uint8 (irb_restore_image in include/irb200.h), for every image (and
every chain stage) of the batch. Default workload is BASELINE.json configs[1] ("cfg2"): Restormer Motion_Deblurring
on a synthetic 1280x720 image, tile 256 / overlap 32, bf16 storage with fp32 accumulation, seeded random-init weights
(no checkpoints ship with the reference).

  value : whole-job MP/s (restored INPUT image pixels, not executed tile pixels) with the uint8 inputs already in HBM
  e2e   : the same through the host API with HOST buffers: pinned H2D of the uint8 images and D2H of the uint8
          results inside the timed region, every step
  N > 1 : cfg2 / cfg5 / cfg1 / swinir_dn — weak scaling: every rank restores its own image per step (independent
          units, no data-path collective), uint8 results gathered to rank 0 over NCCL (asynchronously, overlapped with
          the next step) inside the timed region. cfg3 / cfg4 — the fixed batch is sharded over the ranks (strong).
          Every N > 1 line also carries "strong_4k": ONE 3840x2160 image (cfg5) over the N ranks through
          parallel.restore_strips (local strip blend, seam exchange, uint8 strip gather) next to the same code on
          rank 0 alone.
  roofline     : dominant kernel (by device time) of a per-launch CUDA-event profile of one more step, plus
                 roofline.model_level: the whole step against the model's tensor ceiling and ideal fused traffic
  parity       : one workload tile through the CUDA path vs the oracle (max-abs, PSNR, uint8 LSB histogram), outside
                 the timed region
  cpu_baseline : the oracle (CPU fp32 restatement of the reference, pinned to it by tests/golden) on a bounded
                 sample of the same workload, all host threads; rank 0, N = 1 only
  --impl reference : times that CPU path alone and prints the same JSON line with "impl": "reference"
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "restored megapixels/sec (tiled tool body)"
UNIT = "MP/s"

# A stage = one tool call: (arch, task, tile, overlap, swinir scale, large_model)
# A workload = (stages, (h, w, c), batch images per step, precision, sharding for N > 1, description)
WORKLOADS = {
    "cfg2": dict(stages=[("restormer", "Motion_Deblurring", 256, 32, 1, False)], shape=(720, 1280, 3), batch=1,
                 prec="bf16", multi="weak",
                 desc="cfg2: Restormer Motion_Deblurring 1280x720 synthetic, tile 256 overlap 32 (24 tiles), bf16"),
    "cfg5": dict(stages=[("restormer", "Real_Denoising", 512, 64, 1, False)], shape=(2160, 3840, 3), batch=1,
                 prec="bf16", multi="weak",
                 desc="cfg5: Restormer Real_Denoising 3840x2160 synthetic, tile 512 overlap 64 (45 tiles), bf16"),
    "cfg1": dict(stages=[("swinir", "gray_dn", None, 32, 1, False)], shape=(256, 256, 1), batch=1, prec="fp32",
                 multi="weak",
                 desc="cfg1: SwinIR-M gray_dn 256x256 synthetic, whole image (264x264 after the script pad), fp32"),
    "cfg3": dict(stages=[("swinir", "classical_sr", 256, 32, 4, False)], shape=(512, 512, 3), batch=64, prec="bf16",
                 multi="batch",
                 desc="cfg3: SwinIR-M classical_sr x4 512x512 -> 2048x2048, batch 64, tile 256 overlap 32 "
                      "(9 tiles per image), bf16"),
    "cfg3_whole": dict(stages=[("swinir", "classical_sr", None, 32, 4, False)], shape=(512, 512, 3), batch=64,
                       prec="bf16", multi="batch",
                       desc="cfg3 without tiling (the executor never passes --tile; SURVEY 8d: 'also report whole-image "
                            "mode'): SwinIR-M classical_sr x4 512x512 -> 2048x2048, batch 64, whole 520x520 padded image "
                            "per forward, bf16"),
    "cfg3_1img": dict(stages=[("swinir", "classical_sr", 256, 32, 4, False)], shape=(512, 512, 3), batch=1,
                      prec="bf16", multi="weak",
                      desc="one image of cfg3: SwinIR-M classical_sr x4 512x512 -> 2048x2048, tile 256 overlap 32, bf16"),
    "cfg4": dict(stages=[("restormer", "Deraining", None, 32, 1, False), ("swinir", "color_dn", None, 32, 1, False),
                         ("swinir", "real_sr", None, 32, 4, True)], shape=(512, 512, 3), batch=32, prec="bf16",
                 multi="batch",
                 desc="cfg4: chain Restormer Deraining -> uint8 -> SwinIR-M color_dn -> uint8 -> SwinIR-L real_sr x4 on "
                      "512x512, batch 32, whole-image calls (the executor never passes --tile), uint8 hand-off in HBM, "
                      "bf16"),
    "xrestormer_512": dict(stages=[("xrestormer", "dehazing", None, 0, 1, False)], shape=(512, 512, 3), batch=1,
                           prec="bf16", multi="weak",
                           desc="X-Restormer (dehazing config) 512x512 synthetic, whole image (the BasicSR tool path: no "
                                "tile, pad to 64), bf16"),
    "hat_x4": dict(stages=[("hat", "real_sr_x4", 512, 32, 4, False)], shape=(1024, 1024, 3), batch=1, prec="bf16",
                   multi="weak",
                   desc="HAT (Real_HAT_GAN_SRx4, executor/super_resolution/configs/hat.yml) x4 1024x1024 -> 4096x4096 "
                        "synthetic, HATModel tile mode: tile_size 512 / tile_pad 32 (4 tiles of 544x544, pasted without "
                        "blending), bf16"),
    "swinir_dn_256": dict(stages=[("swinir", "color_dn", None, 32, 1, False)], shape=(256, 256, 3), batch=1,
                          prec="bf16", multi="weak",
                          desc="SwinIR-M color_dn 256x256 synthetic, whole image (the executor's untiled small-image "
                               "call, multitask_tools.py:120-133), bf16"),
    "swinir_dn": dict(stages=[("swinir", "color_dn", None, 32, 1, False)], shape=(512, 512, 3), batch=1, prec="bf16",
                      multi="weak", desc="SwinIR-M color_dn 512x512 synthetic, whole image, bf16"),
}

# BASELINE.md §2 / SURVEY §8(a,d): matmul+conv FLOPs per input pixel of each reference model, and the ideal-fused HBM
# traffic per pixel in bf16 (one read + one write of the residual stream per fusion boundary).
MODEL_FLOPS_PER_PX = {("restormer", 3): 4726624, ("restormer", 1): 4721440, ("swinir", "gray_dn"): 24410160,
                      ("swinir", "color_dn"): 24423120, ("swinir", "color_jpeg_car"): 24034320,
                      ("swinir", "classical_sr"): 26150616, ("swinir", "real_sr_large"): 61339104}
IDEAL_FUSED_BYTES_PER_PX = {"restormer": 18e3, "swinir": 55e3, "swinir_large": 110e3}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tensor=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    tensor_burst=d["bf16_tflops"], source="measured (MEASURED_PEAKS.json; sustained bf16 figure)")
    return dict(hbm=6650.0, tensor=1400.0, tensor_burst=1590.0, source="fallback (B200_PROFILING.md)")


def stage_model(stage):
    """-> (arch, cfg, state_dict) with seeded random-init weights (synth.py seed 0)."""
    from image_restoration_agent_b200 import configs, synth
    arch, task, tile, ov, scale, large = stage
    if arch == "restormer":
        cfg = configs.restormer_config(task)
        return arch, cfg, synth.restormer_state_dict(cfg, seed=0)
    if arch == "xrestormer":
        cfg = configs.xrestormer_config()
        return arch, cfg, synth.xrestormer_state_dict(cfg, seed=0)
    if arch == "hat":
        cfg = configs.hat_config()
        return arch, cfg, synth.hat_state_dict(cfg, seed=0)
    cfg = configs.swinir_config(task, scale=scale, large_model=large)
    return arch, cfg, synth.swinir_state_dict(cfg, seed=0)


def _tile_idx(n, tile, ov):
    return list(range(0, n - tile, tile - ov)) + [n - tile]


def executed_px_per_image(stage, h, w):
    """Pixels the model actually runs on for one h x w input image (script pad + tile over-compute)."""
    arch, task, tile, ov, scale, large = stage
    if arch == "hat":                                                        # hat_model.py:16-75
        hp, wp = (h + 15) // 16 * 16, (w + 15) // 16 * 16
        if tile is None:
            return hp * wp
        span = lambda n: [min(a + tile + ov, n) - max(a - ov, 0) for a in range(0, n, tile)]
        return sum(th * tw for th in span(hp) for tw in span(wp))
    if arch == "swinir":
        ws = 7 if "jpeg" in task else 8
        hp, wp = (h // ws + 1) * ws, (w // ws + 1) * ws                      # inference.py:76-78
    elif arch == "xrestormer":
        hp, wp = (h + 63) // 64 * 64, (w + 63) // 64 * 64                    # xrestormer_model.py:56-66
    else:
        hp, wp = (h + 7) // 8 * 8, (w + 7) // 8 * 8                          # demo.py:122-126
    if tile is None:
        return hp * wp
    t = min(tile, hp, wp)
    return len(_tile_idx(hp, t, ov)) * len(_tile_idx(wp, t, ov)) * t * t


def model_level(wl, ms_step, images_per_step, step_bytes, peaks):
    """The whole step against the model's ceiling (290 / 56 / 52 / 22 MP/s at the sustained bf16 peak,
    or the ideal-fused HBM traffic when that binds) and the bytes it moved against the ideal fused traffic."""
    h, w, c = wl["shape"]
    flops = ideal = px_exec = 0.0
    hh, ww = h, w
    for st in wl["stages"]:
        arch, task, tile, ov, scale, large = st
        key = ("restormer", c) if arch == "restormer" else ("swinir", "real_sr_large" if large else task)
        if arch in ("xrestormer", "hat") or key not in MODEL_FLOPS_PER_PX:
            return {"note": "no model-level ceiling: BASELINE.md section 2 has no FLOP count for this model"}
        px = executed_px_per_image(st, hh, ww)
        px_exec += px
        flops += MODEL_FLOPS_PER_PX[key] * px
        ideal += IDEAL_FUSED_BYTES_PER_PX["swinir_large" if large else arch] * px
        hh, ww = hh * scale, ww * scale
    over = px_exec / (h * w * len(wl["stages"]))          # executed / restored pixels (pad + tile overlap)
    flops *= images_per_step
    ideal *= images_per_step
    sec = ms_step / 1e3
    t_tensor, t_hbm = flops / (peaks["tensor"] * 1e12), ideal / (peaks["hbm"] * 1e9)
    mp = images_per_step * h * w / 1e6
    ceil_exec = mp / max(t_tensor, t_hbm)                 # restored MP/s if every EXECUTED pixel ran at the ceiling
    ceil_untiled = ceil_exec * over                       # BASELINE.md §2's per-model figure (no tile over-compute)
    return {"mp_s_ceiling": round(ceil_untiled, 1), "ceiling_bound": "tensor" if t_tensor >= t_hbm else "hbm",
            "frac_of_ceiling": round((mp / sec) / ceil_untiled, 4),
            "overcompute": round(over, 3), "frac_of_ceiling_executed_px": round((mp / sec) / ceil_exec, 4),
            "model_gflop_executed": round(flops / 1e9, 1),
            "tensor_frac": round(flops / sec / 1e12 / peaks["tensor"], 4),
            "step_bytes": int(step_bytes), "ideal_fused_bytes": int(ideal),
            "traffic_ratio": round(step_bytes / ideal, 2) if ideal else None,
            "note": "mp_s_ceiling = 1 / max(model FLOPs per px / sustained bf16 peak, ideal fused bytes per px / HBM "
                    "peak) (BASELINE.md section 2); frac_of_ceiling = value / mp_s_ceiling counts restored pixels, "
                    "frac_of_ceiling_executed_px credits the padded / overlapping tile pixels the reference loop also "
                    "executes; step_bytes = sum of the per-kernel algorithmic bytes launched in one step"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.lines, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [s for s, p in zip(sm, pw) if p >= 0.5 * max(pw)] if pw else sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}



# ------------------------------------------------------------------------------------------------------------
# Shared by both arms: the config dict (identical in the two JSON lines), the oracle models
# ------------------------------------------------------------------------------------------------------------
def config_dict(name):
    wl = WORKLOADS[name]
    return {"workload": wl["desc"], "weights": "seeded random init (synth.py seed 0)",
            "images_per_step_per_job": wl["batch"],
            "l2": "every step streams a multi-GB activation workspace (>> 126 MB L2) between reuses; no explicit flush",
            "stages": [f"{a}:{t}" + (f":tile{ti}/ov{ov}" if ti else ":whole-image") for a, t, ti, ov, s, lg in wl["stages"]]}


def oracle_model(stage):
    from oracle import hat_ref, restormer_ref, swinir_ref, xrestormer_ref
    arch, cfg, sd = stage_model(stage)
    cls = {"restormer": restormer_ref.Model, "swinir": swinir_ref.Model, "xrestormer": xrestormer_ref.Model,
           "hat": hat_ref.Model}[arch]
    return cls(sd, cfg), cfg


# ------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle on a bounded sample of the workload
# ------------------------------------------------------------------------------------------------------------
def cpu_sample_mps(name: str, steps: int, warmup: int, threads: int | None = None, budget_s: float = 150.0):
    """Per stage, one step = the oracle model on ONE tile of the stage's tile list (the unit of the reference loop,
    demo.py:145-148 / inference.py:312-315), or on a 256x256 crop of a whole-image call. The tile loop, the batch
    loop and the chain are sequential in the reference, so seconds per image = sum over stages of
    (n_tiles x s/tile) and MP/s = image MP / that; the extrapolation is spelled out in `sample`."""
    import numpy as np
    import torch
    from image_restoration_agent_b200 import synth
    from oracle import tiling

    wl = WORKLOADS[name]
    h, w, c = wl["shape"]
    ncores = os.cpu_count() or 1
    torch.set_num_threads(threads or ncores)
    img = synth.synthetic_image_u8(h, w, c, seed=0)
    x0 = torch.from_numpy(img).float().div(255.0).permute(2, 0, 1).unsqueeze(0)
    sec_img, notes, first_ms = 0.0, [], None
    t_begin = time.perf_counter()
    hh, ww = h, w
    per_stage_budget = budget_s / len(wl["stages"])
    for st in wl["stages"]:
        arch, task, tile, ov, scale, large = st
        model, cfg = oracle_model(st)
        x = x0 if (hh, ww) == (h, w) else torch.nn.functional.interpolate(x0, size=(hh, ww), mode="nearest")
        if arch == "swinir":
            x = tiling.swinir_pad(x, cfg["window_size"])
        hp, wp = x.shape[2:]
        if arch == "hat":   # crop-with-pad tiles of several shapes: time a 128x128 crop, scale by the executed pixels
            th, tw, ntiles = 128, 128, 1
            factor = executed_px_per_image(st, hh, ww) / float(th * tw)
        elif tile is None:
            th, tw, ntiles = min(hp, 256), min(wp, 256), 1
            factor = (hp * wp) / float(th * tw)
        else:
            t = min(tile, hp, wp)
            th = tw = t
            ntiles = len(tiling.tile_indices(hp, t, ov)) * len(tiling.tile_indices(wp, t, ov))
            factor = float(ntiles)
        patch = x[..., :th, :tw].contiguous()
        times = []
        t_stage = time.perf_counter()
        with torch.no_grad():
            for i in range(warmup + steps):
                t0 = time.perf_counter()
                y = model(patch)
                dt = time.perf_counter() - t0
                if i >= warmup:
                    times.append(dt)
                if times and time.perf_counter() - t_stage > per_stage_budget:
                    break
        assert np.isfinite(y.numpy()).all()
        sec_tile = sum(times) / len(times)
        first_ms = first_ms if first_ms is not None else sec_tile * 1e3
        sec_img += sec_tile * factor
        notes.append(f"{arch}:{task} {len(times)} timed forward(s) of one {th}x{tw} "
                     f"{'tile' if tile and arch != 'hat' else 'crop'} ({sec_tile:.3f} s each) x {factor:.2f}")
        hh, ww = hh * scale, ww * scale
    mps = (h * w / 1e6) / sec_img
    sample = (f"oracle fp32 on {torch.get_num_threads()} threads; " + "; ".join(notes) +
              f"; MP/s = image MP / sum of (s per unit x units per image); batch of {wl['batch']} = that many "
              f"sequential images in the reference")
    return mps, sec_img * 1e3, torch.get_num_threads(), sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    wl = WORKLOADS[args.workload]
    mps, ms_img, cores, sample = cpu_sample_mps(args.workload, max(1, args.steps), max(0, min(args.warmup, 1)))
    line = {
        "impl": "reference", "metric": METRIC, "value": mps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_img * wl["batch"], "higher_is_better": True,
        "scaling": "strong" if wl["multi"] == "batch" else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_dict(args.workload),
        "arm": "reference algorithm on host CPU cores (oracle port of the PyTorch modules; /root/reference does not "
               "travel to the GPU box)",
        "cpu_baseline": {"value": mps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": mps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def parity_report(eng, stage, shape, prec):
    """One workload tile through the CUDA path (Engine.forward = model(x) behind the C ABI) vs the oracle on the same
    input: max-abs and PSNR on the [0,1] scale, LSB histogram of clamp/rint(x*255). Outside every timed region."""
    import math

    import numpy as np
    import torch
    from image_restoration_agent_b200 import synth
    from oracle import tiling

    arch, task, tile, ov, scale, large = stage
    h, w, c = shape
    model, cfg = oracle_model(stage)
    img = synth.synthetic_image_u8(h, w, c, seed=0)
    x = torch.from_numpy(img).float().div(255.0).permute(2, 0, 1).unsqueeze(0)
    if arch == "swinir":
        x = tiling.swinir_pad(x, cfg["window_size"])
    t = min(tile or (256 if arch == "restormer" else 128), x.shape[2], x.shape[3])
    if arch in ("xrestormer", "hat"):
        t = 128
    if arch == "swinir" and tile is None:
        t = min(t, 128)                                   # 128^2 tokens at full depth: ~3 s of oracle time
    x = x[..., :t, :t].contiguous()
    t0 = time.perf_counter()
    with torch.no_grad():
        want = model(x).numpy()
    sec = time.perf_counter() - t0
    got = eng.forward(x.to(eng.dev)).cpu().numpy()
    err = float(np.abs(got - want).max())
    mse = float(((got.astype(np.float64) - want) ** 2).mean())
    q = lambda a: np.rint(np.clip(a, 0, 1) * 255.0).astype(np.int64)
    d = np.abs(q(got) - q(want))
    hist = np.bincount(d.ravel(), minlength=4)[:8].tolist()
    tol = 1e-4 if prec == "fp32" else 2e-2
    return {"tile": f"{arch}:{task} first {t}x{t} tile of the workload image, full depth", "max_abs": err,
            "psnr_db": round(200.0 if mse == 0 else 10 * math.log10(1.0 / mse), 2), "tolerance": tol,
            "within_tolerance": bool(err <= tol), "uint8_max_lsb": int(d.max()), "uint8_lsb_histogram": hist,
            "frac_within_1_lsb": round(float((d <= 1).mean()), 6), "oracle_seconds": round(sec, 2)}


# ------------------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------------------
class Job:
    """The engines of a workload's stages and one step over a list of device images (the chain keeps the uint8
    hand-off between stages in HBM, T7 semantics)."""

    def __init__(self, name, device, tile_batch=0, graph=False, options=()):
        from image_restoration_agent_b200._lib import check
        from image_restoration_agent_b200.engine import Engine
        self.wl = WORKLOADS[name]
        self.graph = graph
        self.engines = []
        for st in self.wl["stages"]:
            arch, cfg, sd = stage_model(st)
            e = Engine(arch, cfg, sd, precision=self.wl["prec"], device=device)
            if tile_batch:
                e.set_tile_batch(tile_batch)
            for kv in options:   # library options (include/irb200.h: irb_set_option), e.g. fuse_pwdw=0 for an A/B
                k, v = kv.split("=")
                check(e.lib.irb_set_option(e._h, k.encode(), int(v)))
            self.engines.append(e)
        self.out_scale = 1
        for e in self.engines:
            self.out_scale *= e.scale

    def restore(self, img_dev):
        x = img_dev
        for e, st in zip(self.engines, self.wl["stages"]):
            if st[0] == "hat":   # HATModel's crop-with-pad tile mode: st[2] = tile_size, st[3] = tile_pad
                x, _ = e.restore_crop_device(x, st[2], st[3], bgr=True)
            else:
                x, _ = e.restore_device(x, st[2], st[3], bgr=(x.shape[2] == 3), graph=self.graph)
        return x

    def close(self):
        for e in self.engines:
            e.close()


def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from image_restoration_agent_b200 import parallel, synth
    from image_restoration_agent_b200.profiler import KernelProfile

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the irb200 path has no CPU fallback (use --impl reference "
                         "for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    wl = WORKLOADS[args.workload]
    h, w, c = wl["shape"]
    job = Job(args.workload, local, args.tile_batch, graph=args.graph, options=args.opt)
    s = job.out_scale
    batch_mode = wl["multi"] == "batch"
    # images this rank restores per step: its share of the fixed batch (strong), or one own image (weak)
    if batch_mode:
        lo, hi = parallel.partition(wl["batch"], world)[rank]
        seeds = list(range(lo, hi))
        images_per_step = wl["batch"]
    else:
        seeds = [rank]
        images_per_step = world
    imgs_np = [synth.synthetic_image_u8(h, w, c, seed=sd) for sd in seeds]
    imgs_dev = [torch.from_numpy(a).to(dev) for a in imgs_np]
    imgs_pin = [torch.from_numpy(a).pin_memory() for a in imgs_np]
    n_loc = len(seeds)
    out_pin = torch.empty((max(n_loc, 1), h * s, w * s, c), dtype=torch.uint8).pin_memory()
    out_dev = torch.empty((max(n_loc, 1), h * s, w * s, c), dtype=torch.uint8, device=dev)
    mp_step = images_per_step * h * w / 1e6
    ag = parallel.AsyncGather(out_dev, world, rank) if world > 1 else None
    gathered_bytes = out_dev.numel() * (world - 1) if world > 1 else 0

    def step_resident():
        for i, d in enumerate(imgs_dev):
            out_dev[i].copy_(job.restore(d), non_blocking=True)
        if ag is not None:
            ag.submit(out_dev)

    def step_e2e():
        for i, p in enumerate(imgs_pin):
            d = p.to(dev, non_blocking=True)
            out_dev[i].copy_(job.restore(d), non_blocking=True)
        if ag is not None:
            k = ag.submit(out_dev)
            ag.work[k].wait()
            if rank == 0:
                out_pin.copy_(ag.recv[k][0], non_blocking=True)
        else:
            out_pin.copy_(out_dev, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        if ag is not None:
            ag.flush()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(args.warmup, 3)):
        step_resident()
    if ag is not None:
        ag.flush()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    ms_total = timed(step_resident, args.steps)
    clocks = sampler.stop() if sampler else None
    ms_step = ms_total / args.steps
    value = mp_step / (ms_step / 1e3)

    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps) / args.steps
    e2e_value = mp_step / (ms_e2e / 1e3)

    # ---- per-launch profile of one more image (CUDA events around every kernel, on the launching stream) ----
    torch.cuda.synchronize()
    job_prof = job
    if args.graph:   # launches inside a replayed graph cannot be bracketed: profile the eager path of the same code
        job_prof = Job(args.workload, local, args.tile_batch, graph=False, options=args.opt)
        job_prof.restore(imgs_dev[0])   # sizes its workspace outside the profiled call
        torch.cuda.synchronize()
    with KernelProfile() as prof:
        job_prof.restore(imgs_dev[0] if imgs_dev else torch.from_numpy(synth.synthetic_image_u8(h, w, c, 0)).to(dev))
    launches_per_image = prof.launches
    launches_per_step = launches_per_image * n_loc

    strong4k = None
    if args.strong4k and (world > 1 or args.strong4k > 1):
        strong4k = strong_4k(job if args.workload == "cfg5" and not args.graph else None, dev, world, rank, local)

    if rank == 0:
        peaks = load_peaks()
        ridge = peaks["tensor"] * 1e12 / (peaks["hbm"] * 1e9)
        table = []
        for e in prof.entries:
            sec = e["ms"] / 1e3
            ai = e["flops"] / max(e["bytes"], 1.0)
            bound = "tensor" if ai > ridge else "hbm"
            ach = e["flops"] / sec / 1e12 if bound == "tensor" else e["bytes"] / sec / 1e9
            peak = peaks["tensor"] if bound == "tensor" else peaks["hbm"]
            table.append({"kernel": e["name"], "launches": e["launches"], "ms": round(e["ms"], 4),
                          "share": round(e["ms"] / prof.total_ms, 4), "bound": bound, "achieved": round(ach, 2),
                          "unit": "TFLOP/s" if bound == "tensor" else "GB/s", "frac": round(ach / peak, 4),
                          "gflop": round(e["flops"] / 1e9, 3), "mbytes": round(e["bytes"] / 1e6, 3)})
        top = table[0]
        top_e = prof.entries[0]
        roofline = {
            "kernel": top["kernel"], "bound": top["bound"], "achieved": top["achieved"],
            "peak": peaks["tensor"] if top["bound"] == "tensor" else peaks["hbm"], "unit": top["unit"],
            "frac": top["frac"], "traffic": None, "launches_per_image": top["launches"],
            "avg_launch_ms": round(top_e["ms"] / top_e["launches"], 5), "share_of_step": top["share"],
            "algorithmic_per_launch": {"flops": top_e["flops"] / top_e["launches"],
                                       "bytes": top_e["bytes"] / top_e["launches"]},
            "peak_source": peaks["source"],
            "how": "CUDA events recorded by the library around every launch of one extra image after the timed region",
        }
        # DRAM traffic of the dominant kernel: profiles/traffic.json holds, per kernel class, the ratio
        # (dram__bytes_read.sum + dram__bytes_write.sum) / algorithmic bytes measured by one ncu --set full capture of a
        # launch of that class (file and launch named there); traffic per launch = ratio x algorithmic bytes per launch.
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            tj = json.load(open(tpath)).get(top["kernel"])
            if tj and top["bound"] == "hbm":
                roofline["traffic"] = round(tj["dram_over_algorithmic"] * roofline["algorithmic_per_launch"]["bytes"])
                roofline["traffic_source"] = tj["source"]
        tot_flops = sum(e["flops"] for e in prof.entries)
        tot_bytes = sum(e["bytes"] for e in prof.entries)
        per_rank_images = max(n_loc, 1)
        ms_image = ms_step / per_rank_images
        roofline["model_level"] = model_level(wl, ms_image, 1, tot_bytes, peaks)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong" if batch_mode else "weak",
            "vs_baseline": None, "dtype": wl["prec"] if wl["prec"] != "fp32" else "f32", "data": "synthetic",
            "config": config_dict(args.workload),
            "run": {"images_per_step": images_per_step, "images_per_step_this_rank": n_loc,
                    "tile_batch": args.tile_batch or "auto", "cuda_graph_replay": bool(args.graph),
                    "library_options": list(args.opt),
                    "workspace_gib": round(sum(e._ws.numel() for e in job.engines if e._ws is not None) / 2**30, 2),
                    "gathered_bytes_per_step": gathered_bytes,
                    "gather": "async NCCL gather of the uint8 results to rank 0, overlapped with the next step, flushed "
                              "before the closing timestamp" if world > 1 else None},
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(sum(a.nbytes for a in imgs_np)),
                    "d2h_bytes_per_step": int(out_pin.numel() if n_loc else 0)},
            "gpu_launches": launches_per_step * args.steps,
            "roofline": roofline,
            "step_totals": {"device_ms_profiled_one_image": round(prof.total_ms, 3),
                            "algorithmic_gflop_one_image": round(tot_flops / 1e9, 1),
                            "algorithmic_gbytes_one_image": round(tot_bytes / 1e9, 2),
                            "tensor_frac_of_step": round(tot_flops / (ms_image / 1e3) / 1e12 / peaks["tensor"], 4),
                            "hbm_frac_of_step": round(tot_bytes / (ms_image / 1e3) / 1e9 / peaks["hbm"], 4)},
            "clocks": clocks,
            "kernels": table[:args.top],
        }
        if strong4k is not None:
            line["strong_4k"] = strong4k
        if world == 1 and not args.no_cpu:
            line["parity"] = [parity_report(e, st, wl["shape"] if k == 0 else
                                            (wl["shape"][0], wl["shape"][1], wl["shape"][2]), wl["prec"])
                              for k, (e, st) in enumerate(zip(job_prof.engines, wl["stages"]))]
            mps, ms_img, cores, sample = cpu_sample_mps(args.workload, args.cpu_steps, 1,
                                                        budget_s=60.0 if len(wl["stages"]) == 1 else 120.0)
            line["cpu_baseline"] = {"value": mps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if job_prof is not job:
        job_prof.close()
    job.close()
    return 0


def strong_4k(job, dev, world, rank, local, warmup=2, steps=5):
    """north_star's multi-GPU path on BASELINE config 5: ONE 3840x2160 image, its 45 tiles sharded over the N ranks
    (parallel.restore_strips: local strip blend, seam-row exchange, uint8 strip gather to rank 0), timed on the device
    as the max over ranks, next to the SAME code on rank 0 alone (solo=True) in the same process and clock state."""
    import torch
    import torch.distributed as dist
    from image_restoration_agent_b200 import parallel, synth

    own = job is None
    if own:
        job = Job("cfg5", local)
    eng = job.engines[0]
    wl = WORKLOADS["cfg5"]
    h, w, c = wl["shape"]
    tile, ov = wl["stages"][0][2], wl["stages"][0][3]
    img = torch.from_numpy(synth.synthetic_image_u8(h, w, c, seed=0)).to(dev)

    def timed(fn, k, all_ranks):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) if (all_ranks or rank == 0) else 0.0], device=dev)
        if world > 1:
            dist.barrier()
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / k

    stats = {}

    def step_n():
        nonlocal stats
        _, stats = parallel.restore_strips(eng, img, tile, ov)

    def step_1():
        if rank == 0:
            parallel.restore_strips(eng, img, tile, ov, solo=True)

    for _ in range(warmup):
        step_n()
    ms_n = timed(step_n, steps, True)
    for _ in range(warmup):
        step_1()
    ms_1 = timed(step_1, steps, False)
    # bit-identity of the two results, checked once outside the timed regions
    out_n, _ = parallel.restore_strips(eng, img, tile, ov)
    same = None
    if rank == 0:
        out_1, _ = parallel.restore_strips(eng, img, tile, ov, solo=True)
        same = bool(torch.equal(out_n, out_1))
    seam = torch.tensor([float(stats.get("seam_bytes", 0))], device=dev)
    if world > 1:
        dist.all_reduce(seam, op=dist.ReduceOp.SUM)
    if own:
        job.close()
    mp = h * w / 1e6
    nt = 45
    per_rank = -(-nt // world)
    return {"workload": wl["desc"] + "; ONE image over all ranks (parallel.restore_strips)", "n_gpus": world,
            "value": mp / (ms_n / 1e3), "unit": UNIT, "ms_per_image": ms_n,
            "single_gpu_value": mp / (ms_1 / 1e3), "single_gpu_ms_per_image": ms_1,
            "speedup_over_1_gpu": ms_1 / ms_n, "ceiling_from_tile_imbalance": round(nt / per_rank, 2),
            "seam_bytes_all_ranks": int(seam.item()), "gathered_strip_bytes": int(stats.get("gather_bytes", 0)),
            "bit_identical_to_single_gpu": same, "steps": steps, "warmup": warmup,
            "timing": "CUDA events, max over ranks, barrier + synchronize on both sides; input image resident in HBM "
                      "on every rank, result uint8 on rank 0"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=0, help="timed steps (default: 10, or 2 for the batch workloads)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--tile-batch", type=int, default=0)
    ap.add_argument("--graph", type=int, default=-1,
                    help="1: replay a CUDA graph of the whole tool body per image (Engine.restore_device(graph=True)); "
                         "0: issue the launches from host code every step; default: per workload")
    ap.add_argument("--opt", action="append", default=[], help="library option key=value (irb_set_option), repeatable")
    ap.add_argument("--cpu-steps", type=int, default=3, help="timed oracle tile forwards of the cpu_baseline leg")
    ap.add_argument("--strong4k", type=int, default=1,
                    help="N > 1: append the strong_4k object (one 4K image over the N ranks); 0 skips it; 2 also runs "
                         "it at N = 1")
    ap.add_argument("--no-cpu", action="store_true", help="skip the parity and cpu_baseline legs (ncu / quick runs)")
    ap.add_argument("--top", type=int, default=12, help="kernel table rows in the JSON line")
    args = ap.parse_args()
    if args.steps <= 0:
        args.steps = 2 if WORKLOADS[args.workload]["multi"] == "batch" else 10
    if args.graph < 0:
        args.graph = GRAPH_DEFAULT.get(args.workload, 0)
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


# CUDA-graph replay per workload (decided by the A/B in profiles/: replay wins where the step is launch-bound)
GRAPH_DEFAULT = {}


if __name__ == "__main__":
    sys.exit(main())
