#!/bin/bash
# ncu --set full of the four GEMMs of one full-resolution Restormer block (qkv, attn.v, project_in, project_out), unfused FFN
mkdir -p gpurun_out
export IRB_FUSED_FFN=0
CMD="python bench.py --steps 1 --warmup 0 --no-cpu"
timeout 300 $CMD > gpurun_out/ncu_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:tc_gemm --launch-skip ${2:-146} --launch-count 4 -f -o gpurun_out/$1 $CMD > gpurun_out/ncu_$1.log 2>&1
ls -la gpurun_out/$1.ncu-rep
