#!/bin/bash
# quick iteration on the GEMM epilogue: GPU tests, then cfg2 bench (unfused FFN) with the specialised epilogue /
# TMA-staged residual switched on and off. usage: gpu_iter4.sh TAG
TAG=${1:-it}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/${TAG}_pytest.log
export IRB_FUSED_FFN=0
run() { # name env...
  local n=$1; shift
  env "$@" timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/${TAG}_$n.json 2> gpurun_out/${TAG}_$n.err; echo "$n rc=$?"
  python tools/show_bench.py gpurun_out/${TAG}_$n.json 2>/dev/null | grep -E "^==|tc_gemm|tc_conv" | head -${TOPN:-14}
}
run spec IRB_GEMM_SPEC=1
run spec_noresTMA IRB_GEMM_RESTMA=0
run generic IRB_GEMM_SPEC=0 IRB_GEMM_RESTMA=0
