#!/bin/bash
# tests + bench (cfg2, swinir_dn) + ncu full captures of the decoder-level kernels. usage: gpu_iter2.sh TAG
set -x
TAG=${1:-it}
bash tools/gpu_iter.sh "" cfg2 swinir_dn
grep -q "rc=0" gpurun_out/pytest_iter.log || { tail -30 gpurun_out/pytest_iter.log; exit 1; }
bash tools/gpu_ncu_cfg2.sh $TAG
