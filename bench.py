#!/usr/bin/env python
"""bench.py — restored megapixels/s of the tiled restoration tool body on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg5|cfg1|cfg3] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one pass of the hot path over one image of the workload: pad -> tiles -> model forward on every tile
-> overlap blend -> crop -> clamp -> round -> uint8 (irb_restore_image in include/irb200.h). Default workload is
BASELINE.json configs[1] ("cfg2"): Restormer Motion_Deblurring on a synthetic 1280x720 image, tile 256 / overlap
32, bf16 storage with fp32 accumulation, seeded random-init weights (no checkpoints ship with the reference).

  value : whole-job MP/s (restored image pixels, not executed tile pixels) with the uint8 input already in HBM
  e2e   : the same through Engine.restore() with HOST buffers: pinned H2D of the uint8 image and D2H of the
          uint8 result inside the timed region, every step
  N > 1 : weak scaling — every rank restores its own image per step (independent units, no data-path collective),
          the uint8 results are gathered to rank 0 over NCCL inside the timed region; time = max over ranks
  roofline     : dominant kernel (by device time) of a per-launch CUDA-event profile of one more step
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

# name -> (arch, task, (h, w, c), tile, overlap, precision, description)
WORKLOADS = {
    "cfg2": ("restormer", "Motion_Deblurring", (720, 1280, 3), 256, 32, "bf16",
             "cfg2: Restormer Motion_Deblurring 1280x720 synthetic, tile 256 overlap 32 (24 tiles), bf16"),
    "cfg5": ("restormer", "Real_Denoising", (2160, 3840, 3), 512, 64, "bf16",
             "cfg5: Restormer Real_Denoising 3840x2160 synthetic, tile 512 overlap 64 (45 tiles), bf16"),
    "cfg1": ("swinir", "gray_dn", (256, 256, 1), None, 32, "fp32",
             "cfg1: SwinIR-M gray_dn 256x256 synthetic, whole image (264x264 after the script pad), fp32"),
    "cfg3": ("swinir", "classical_sr", (512, 512, 3), 256, 32, "bf16",
             "cfg3 (one image of the batch): SwinIR-M classical_sr x4 512x512 -> 2048x2048, tile 256 overlap 32, bf16"),
    "swinir_dn": ("swinir", "color_dn", (512, 512, 3), None, 32, "bf16",
                  "SwinIR-M color_dn 512x512 synthetic, whole image, bf16"),
}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tensor=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    tensor_burst=d["bf16_tflops"], source="measured (MEASURED_PEAKS.json; sustained bf16 figure)")
    return dict(hbm=6650.0, tensor=1400.0, tensor_burst=1590.0, source="fallback (B200_PROFILING.md)")


def build_model(name):
    from image_restoration_agent_b200 import configs, synth
    arch, task, shape, tile, ov, prec, desc = WORKLOADS[name]
    if arch == "restormer":
        cfg = configs.restormer_config(task)
        sd = synth.restormer_state_dict(cfg, seed=0)
    else:
        cfg = configs.swinir_config(task, scale=4 if task == "classical_sr" else 1)
        sd = synth.swinir_state_dict(cfg, seed=0)
    return arch, cfg, sd


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
# CPU arm: the oracle on a bounded sample of the workload
# ------------------------------------------------------------------------------------------------------------
def cpu_sample_mps(name: str, steps: int, warmup: int, threads: int | None = None, budget_s: float = 150.0):
    """One step = the oracle model on ONE tile of the workload's tile list (the unit of the reference loop,
    demo.py:145-148 / inference.py:312-315). MP/s = image pixels / (n_tiles * mean seconds per tile): the tile
    loop is sequential in the reference, so the per-tile time extrapolates linearly (labelled in `sample`)."""
    import numpy as np
    import torch
    from image_restoration_agent_b200 import synth
    from oracle import restormer_ref, swinir_ref, tiling

    arch, task, (h, w, c), tile, ov, prec, desc = WORKLOADS[name]
    _, cfg, sd = build_model(name)
    ncores = os.cpu_count() or 1
    torch.set_num_threads(threads or ncores)
    model = (restormer_ref.Model if arch == "restormer" else swinir_ref.Model)(sd, cfg)
    img = synth.synthetic_image_u8(h, w, c, seed=0)
    x = torch.from_numpy(img).float().div(255.0).permute(2, 0, 1).unsqueeze(0)
    if arch == "swinir":
        x = tiling.swinir_pad(x, cfg["window_size"])
    hp, wp = x.shape[2:]
    if tile is None:
        th, tw, ntiles = hp, wp, 1
        # a whole-image forward is one unit; bound it by a centre crop when it is large
        if hp * wp > 300 * 300:
            th = tw = 256
    else:
        t = min(tile, hp, wp)
        th = tw = t
        ntiles = len(tiling.tile_indices(hp, t, ov)) * len(tiling.tile_indices(wp, t, ov))
    patch = x[..., :th, :tw].contiguous()
    times = []
    t_begin = time.perf_counter()
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            y = model(patch)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
            if times and time.perf_counter() - t_begin > budget_s:  # keep the whole run within a few minutes
                break
    assert np.isfinite(y.numpy()).all()
    sec_tile = sum(times) / len(times)
    if tile is None:
        sec_img = sec_tile * (hp * wp) / float(th * tw)
    else:
        sec_img = sec_tile * ntiles
    mps = (h * w / 1e6) / sec_img
    sample = (f"oracle fp32 on {torch.get_num_threads()} threads: {len(times)} timed forward(s) of one {th}x{tw} tile "
              f"({sec_tile:.3f} s each); MP/s = image MP / ({ntiles} tiles x s/tile)"
              if tile is not None else
              f"oracle fp32 on {torch.get_num_threads()} threads: {len(times)} timed forward(s) of a {th}x{tw} crop "
              f"({sec_tile:.3f} s each), scaled by padded-image/crop pixels")
    return mps, sec_tile * 1e3, torch.get_num_threads(), sample


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    arch, task, (h, w, c), tile, ov, prec, desc = WORKLOADS[args.workload]
    mps, ms_step, cores, sample = cpu_sample_mps(args.workload, max(1, args.steps), max(0, min(args.warmup, 1)))
    line = {
        "impl": "reference", "metric": METRIC, "value": mps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "weights": "seeded random init (synth.py seed 0)",
                   "arm": "reference algorithm on host CPU cores (oracle port of the PyTorch modules)"},
        "cpu_baseline": {"value": mps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": mps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------------------
def run_b200(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    from image_restoration_agent_b200 import synth
    from image_restoration_agent_b200.engine import Engine
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
    arch, task, (h, w, c), tile, ov, prec, desc = WORKLOADS[args.workload]
    _, cfg, sd = build_model(args.workload)
    eng = Engine(arch, cfg, sd, precision=prec, device=local)
    if args.tile_batch:
        eng.set_tile_batch(args.tile_batch)
    s = eng.scale
    img_np = synth.synthetic_image_u8(h, w, c, seed=rank)
    img_dev = torch.from_numpy(img_np).to(dev)
    img_pin = torch.from_numpy(img_np).pin_memory()
    out_pin = torch.empty((h * s, w * s, c), dtype=torch.uint8).pin_memory()
    gather_buf = [torch.empty((h * s, w * s, c), dtype=torch.uint8, device=dev) for _ in range(world)] \
        if (world > 1 and rank == 0) else None
    mp_img = h * w / 1e6

    strong = bool(getattr(args, "strong", False)) and world > 1
    if strong:  # every rank works on the SAME image (rank 0's)
        from image_restoration_agent_b200 import parallel
        img_np = synth.synthetic_image_u8(h, w, c, seed=0)
        img_dev = torch.from_numpy(img_np).to(dev)
        img_pin = torch.from_numpy(img_np).pin_memory()

    def step_resident():
        if strong:
            u8, _ = parallel.restore_sharded(eng, img_dev, tile, ov)
            return u8
        u8, _ = eng.restore_device(img_dev, tile, ov)
        if world > 1:
            dist.gather(u8, gather_buf, dst=0)
        return u8

    def step_e2e():
        d = img_pin.to(dev, non_blocking=True)
        if strong:
            u8, _ = parallel.restore_sharded(eng, d, tile, ov)
            if rank == 0:
                out_pin.copy_(u8, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            return
        u8, _ = eng.restore_device(d, tile, ov)
        if world > 1:
            dist.gather(u8, gather_buf, dst=0)
            if rank == 0:
                out_pin.copy_(gather_buf[0], non_blocking=True)
        else:
            out_pin.copy_(u8, non_blocking=True)
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
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(args.warmup, 3)):
        step_resident()
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    ms_total = timed(step_resident, args.steps)
    clocks = sampler.stop() if sampler else None
    ms_step = ms_total / args.steps
    value = (1 if strong else world) * mp_img / (ms_step / 1e3)

    for _ in range(2):
        step_e2e()
    ms_e2e = timed(step_e2e, args.steps) / args.steps
    e2e_value = (1 if strong else world) * mp_img / (ms_e2e / 1e3)

    # ---- per-launch profile of one more step (CUDA events around every kernel, on the launching stream) ----
    torch.cuda.synchronize()
    with KernelProfile() as prof:
        eng.restore_device(img_dev, tile, ov)
    launches_per_step = prof.launches
    need_ws = eng._ws.numel() if eng._ws is not None else 0

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
            "frac": top["frac"], "traffic": None, "launches_per_step": top["launches"],
            "avg_launch_ms": round(top_e["ms"] / top_e["launches"], 5), "share_of_step": top["share"],
            "algorithmic_per_launch": {"flops": top_e["flops"] / top_e["launches"],
                                       "bytes": top_e["bytes"] / top_e["launches"]},
            "peak_source": peaks["source"],
            "how": "CUDA events recorded by the library around every launch of one extra step after the timed region",
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
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong" if strong else "weak",
            "vs_baseline": None, "dtype": prec if prec != "fp32" else "f32", "data": "synthetic",
            "config": {"workload": desc, "weights": "seeded random init (synth.py seed 0)",
                       "images_per_step": 1 if strong else world, "tile_batch": args.tile_batch or "auto",
                       "l2": f"each step streams a {need_ws / 2**30:.1f} GiB activation workspace (>> 126 MB L2); "
                             "no explicit flush"},
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": int(img_np.nbytes),
                    "d2h_bytes_per_step": int(out_pin.numel())},
            "gpu_launches": launches_per_step * args.steps,
            "roofline": roofline,
            "step_totals": {"device_ms_profiled": round(prof.total_ms, 3), "algorithmic_gflop": round(tot_flops / 1e9, 1),
                            "algorithmic_gbytes": round(tot_bytes / 1e9, 2),
                            "tensor_frac_of_step": round(tot_flops / (ms_step / 1e3) / 1e12 / peaks["tensor"], 4),
                            "hbm_frac_of_step": round(tot_bytes / (ms_step / 1e3) / 1e9 / peaks["hbm"], 4)},
            "clocks": clocks,
            "kernels": table[:args.top],
        }
        if world == 1 and not args.no_cpu:
            mps, ms_tile, cores, sample = cpu_sample_mps(args.workload, args.cpu_steps, 1)
            line["cpu_baseline"] = {"value": mps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    eng.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS))
    ap.add_argument("--tile-batch", type=int, default=0)
    ap.add_argument("--cpu-steps", type=int, default=3, help="timed oracle tile forwards of the cpu_baseline leg")
    ap.add_argument("--strong", action="store_true",
                    help="N > 1: strong scaling instead of weak — the tiles of ONE image are sharded across the ranks "
                         "(parallel.restore_sharded), fp32 tiles gathered to rank 0 over NCCL, rank 0 blends")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (ncu / quick runs)")
    ap.add_argument("--top", type=int, default=12, help="kernel table rows in the JSON line")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_b200(args)


if __name__ == "__main__":
    sys.exit(main())
