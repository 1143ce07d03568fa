#!/bin/bash
# A/B of IRB_GEMM_COMPACT: micro on the SwinIR shapes, parity tests, cfg3 / swinir_dn / cfg2 bench
export PYTHONPATH=$PWD
mkdir -p gpurun_out
for kn in "180 360" "360 180"; do
  for c in 0 1; do echo "== $kn compact=$c"; IRB_GEMM_COMPACT=$c timeout 120 python tools/gemm_micro.py $kn 2>&1 | tail -2; done
done
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for wl in cfg3 swinir_dn cfg2; do
  for c in 0 1; do
    IRB_GEMM_COMPACT=$c timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/s28c_${wl}_c$c.json 2> gpurun_out/s28c_${wl}_c$c.err; echo "$wl compact=$c rc=$?"
    python tools/show_bench.py gpurun_out/s28c_${wl}_c$c.json 2>/dev/null | grep -E "^==|tc_gemm" | head -6
  done
done
