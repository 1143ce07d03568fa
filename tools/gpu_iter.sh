#!/bin/bash
# Iteration pass: selected op tests, model parity, short bench.  usage: gpu_iter.sh "<pytest -k expr>" [workloads...]
set -x
mkdir -p gpurun_out
K="$1"; shift
timeout 900 python -m pytest tests -m gpu -x -q -k "$K" > gpurun_out/pytest_iter.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_iter.log
tail -15 gpurun_out/pytest_iter.log
for wl in "$@"; do
  timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/iter_$wl.json 2> gpurun_out/iter_$wl.err; echo "bench $wl rc=$?"
  tail -3 gpurun_out/iter_$wl.err
done
