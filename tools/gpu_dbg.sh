#!/bin/bash
# diagnosis: cfg2 bench (unfused FFN) with parts of the GEMM epilogue disabled (results are wrong on purpose)
mkdir -p gpurun_out
export IRB_FUSED_FFN=0
for d in ${DBGS:-0 1 2 4 7}; do
  IRB_GEMM_DBG=$d timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --top 40 > gpurun_out/dbg_$d.json 2> gpurun_out/dbg_$d.err
  echo "== dbg $d"; python tools/show_bench.py gpurun_out/dbg_$d.json 2>/dev/null | grep -E "^==|tc_gemm\[96|tc_gemm\[256,96"
done
