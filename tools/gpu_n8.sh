#!/bin/bash
# Multi-GPU evidence run (gpurun --gpus N): bench lines at N ranks for the listed workloads. usage: gpu_n8.sh TAG N wl...
TAG=$1; N=$2; shift 2
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${TAG}_gpu.txt 2>&1
for wl in "$@"; do
  extra=""; [ $wl != cfg2 ] && extra="--strong4k 0"
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --workload $wl --warmup 3 --top 6 $extra > gpurun_out/${TAG}_bench_${wl}_n$N.json 2> gpurun_out/${TAG}_bench_${wl}_n$N.err; echo "n$N $wl rc=$?"
done
echo done
