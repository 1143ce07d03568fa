#!/bin/bash
# micro + gemm/op parity + bench of the current build. usage: gpu_iter6.sh TAG
TAG=$1
mkdir -p gpurun_out
export PYTHONPATH=$PWD
for kn in "96 96" "256 96" "96 256"; do
  timeout 120 python tools/gemm_micro.py $kn 2>&1 | tail -4
  IRB_GEMM_DBG=224 timeout 120 python tools/gemm_micro.py $kn 2>&1 | tail -1
done
timeout 900 python -m pytest tests/test_gpu_ops.py -x -q -m gpu 2>&1 | tail -3
export IRB_FUSED_FFN=0
for wl in cfg2 cfg5 cfg3; do
  timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/${TAG}_$wl.json 2> gpurun_out/${TAG}_$wl.err; echo "$wl rc=$?"
  python tools/show_bench.py gpurun_out/${TAG}_$wl.json 2>/dev/null | grep -E "^==|tc_gemm" | head -8
done
