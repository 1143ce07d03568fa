#!/bin/bash
# A/B of two library builds on one workload. usage: gpu_ab_so2.sh WORKLOAD PATTERN
export PYTHONPATH=$PWD
mkdir -p gpurun_out
for v in A B A B; do
  cp ab/libirb200_$v.so image_restoration_agent_b200/libirb200.so
  echo "=== build $v"
  timeout 600 python bench.py --workload $1 --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/ab_${v}_$1.json 2> gpurun_out/ab_${v}_$1.err
  python tools/show_bench.py gpurun_out/ab_${v}_$1.json 2>/dev/null | grep -E "^==|$2" | head -4
done
