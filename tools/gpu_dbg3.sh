#!/bin/bash
# gemm_micro over the IRB_GEMM_DBG switches for one shape. usage: gpu_dbg3.sh "K N" [dbg values...]
export PYTHONPATH=$PWD
KN=$1; shift
for d in "$@"; do
  echo "== dbg=$d"
  IRB_GEMM_DBG=$d timeout 120 python tools/gemm_micro.py $KN 2>&1 | tail -3
done
