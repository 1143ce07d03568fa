#!/usr/bin/env python
"""Launch-overhead / per-tile cost of tc_gemm: times ops.gemm1x1 at M = 148 * 128 * t rows for several t (tiles per CTA)
and fits time = F + c * t. usage: gemm_micro.py [K N]"""
import sys, torch
from image_restoration_agent_b200 import ops
K = int(sys.argv[1]) if len(sys.argv) > 1 else 96
N = int(sys.argv[2]) if len(sys.argv) > 2 else 96
KP = (K + 63) // 64 * 64  # padded row strides (the C = 180 streams of SwinIR use 192)
w = (torch.randn(N, KP, device="cuda") / K ** 0.5).bfloat16()[:, :K]
res = []
for t in (1, 2, 4, 8, 16, 32, 64):
    M = 148 * 128 * t
    x = torch.randn(M, (K + 15) // 16 * 16, device="cuda").bfloat16()[:, :K]
    for _ in range(3):
        ops.gemm1x1(x, w)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    e0.record()
    for _ in range(reps):
        ops.gemm1x1(x, w)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1000 / reps
    res.append((t, us))
    print(f"K={K} N={N} tiles/CTA={t:3d} M={M:8d}  {us:8.2f} us/launch  {M * (K + N) * 2 / us / 1e3:8.1f} GB/s")
(t0, u0), (t1, u1) = res[0], res[-1]
c = (u1 - u0) / (t1 - t0)
print(f"fit: per-tile {c:.3f} us, fixed {u0 - c * t0:.2f} us")
