#!/bin/bash
# one ncu --set full capture inside the cfg2 bench command. usage: gpu_ncu_one.sh TAG KERNEL_REGEX SKIP COUNT [workload]
set -x
mkdir -p gpurun_out
WL=${5:-cfg2}
CMD="python bench.py --workload $WL --steps 1 --warmup 0 --no-cpu"
timeout 300 $CMD > gpurun_out/ncu_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:$2 --launch-skip $3 --launch-count $4 -f -o gpurun_out/$1 $CMD > gpurun_out/ncu_$1.log 2>&1
ls -la gpurun_out/$1.ncu-rep
