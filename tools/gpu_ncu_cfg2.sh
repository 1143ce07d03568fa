#!/bin/bash
# ncu --set full captures of the Restormer decoder-level-1 kernels (C=96, 24 tiles of 256x256 = 1.57 M pixels) inside
# the cfg2 bench command. Launch indices: see DESIGN.md (block order qkv-gemm, dw, gram, softmax, attn-v gemm, stats,
# pin gemm, dw-gate, pout gemm; decoder level 1 starts at tc_gemm launch 152 / block 36).
set -x
mkdir -p gpurun_out
TAG=${1:-r01}
CMD="python bench.py --steps 1 --warmup 0 --no-cpu"
NCU="ncu --set full --import-source on --clock-control none"
timeout 300 $CMD > gpurun_out/ncu_plain.log 2>&1 || exit 1
timeout 600 $NCU -k regex:tc_gemm_kernel --launch-skip 153 --launch-count 4 -f -o gpurun_out/${TAG}_tc_gemm $CMD > gpurun_out/ncu_tc.log 2>&1
timeout 600 $NCU -k regex:dw3x3_tma --launch-skip 72 --launch-count 2 -f -o gpurun_out/${TAG}_dw $CMD > gpurun_out/ncu_dw.log 2>&1
timeout 600 $NCU -k regex:row_stats --launch-skip 72 --launch-count 1 -f -o gpurun_out/${TAG}_row_stats $CMD > gpurun_out/ncu_rs.log 2>&1
timeout 600 $NCU -k regex:mdta --launch-skip 108 --launch-count 3 -f -o gpurun_out/${TAG}_mdta $CMD > gpurun_out/ncu_mdta.log 2>&1
ls -la gpurun_out/*.ncu-rep
