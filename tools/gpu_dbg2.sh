#!/bin/bash
# GEMM diagnosis switches on any workload. usage: WL=swinir_dn DBGS="0 32 96 224" PAT="180,576|180,360" gpu_dbg2.sh
mkdir -p gpurun_out
for d in ${DBGS:-0 32 96 224}; do
  IRB_GEMM_DBG=$d timeout 300 python bench.py --workload ${WL:-cfg2} --steps 3 --warmup 3 --no-cpu --top 40 > gpurun_out/dbg2_$d.json 2> gpurun_out/dbg2_$d.err
  echo "== dbg $d"; python tools/show_bench.py gpurun_out/dbg2_$d.json 2>/dev/null | grep -E "${PAT:-tc_gemm}" | head -8
done
