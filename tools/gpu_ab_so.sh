#!/bin/bash
# A/B of two builds of the library inside one call (ab/libirb200_A.so = before, _B.so = after). usage: gpu_ab_so.sh
export PYTHONPATH=$PWD
mkdir -p gpurun_out
for v in A B A B; do
  cp ab/libirb200_$v.so image_restoration_agent_b200/libirb200.so
  echo "=== build $v"
  for kn in "96 96" "256 96" "96 512"; do timeout 120 python tools/gemm_micro.py $kn 2>&1 | tail -2 | head -1; done
  for wl in cfg2 cfg3; do
    timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/ab_${v}_$wl.json 2> gpurun_out/ab_${v}_$wl.err
    python tools/show_bench.py gpurun_out/ab_${v}_$wl.json 2>/dev/null | grep -E "^==|tc_gemm" | head -6
  done
done
cp ab/libirb200_B.so image_restoration_agent_b200/libirb200.so
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
