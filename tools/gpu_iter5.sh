#!/bin/bash
# A/B of an environment switch on the cfg2 bench (unfused FFN). usage: gpu_iter5.sh TAG "name:ENV=V,ENV2=V" ...
TAG=$1; shift
mkdir -p gpurun_out
export IRB_FUSED_FFN=${IRB_FUSED_FFN:-0}
for spec in "$@"; do
  n=${spec%%:*}; envs=${spec#*:}
  env $(echo $envs | tr ',' ' ') timeout 600 python bench.py --workload ${WL:-cfg2} --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/${TAG}_$n.json 2> gpurun_out/${TAG}_$n.err; echo "$n rc=$?"
  python tools/show_bench.py gpurun_out/${TAG}_$n.json 2>/dev/null | grep -E "${PAT:-^==|tc_gemm|gram|dw3x3}" | head -${TOPN:-14}
done
