#!/bin/bash
# Fused-GDFN iteration: op parity, model parity, fused bench, then one ncu --set full capture. usage: gpu_ffn2.sh TAG [ncu]
TAG=${1:-ffn}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k "gdfn_fused" > gpurun_out/${TAG}_op.log 2>&1; echo "op rc=$?"; tail -3 gpurun_out/${TAG}_op.log
timeout 900 python -m pytest tests/test_gpu_models.py -m gpu -x -q > gpurun_out/${TAG}_models.log 2>&1; echo "models rc=$?"; tail -3 gpurun_out/${TAG}_models.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/${TAG}_bench_fused.json 2> gpurun_out/${TAG}_bench_fused.err; echo "bench1 rc=$?"
python tools/show_bench.py gpurun_out/${TAG}_bench_fused.json 2>/dev/null | grep -E "^==|step:|gdfn|dw3x3_gate"
if [ "$2" = "ncu" ]; then
  timeout 900 ncu --set full --import-source on --clock-control none -k regex:gdfn_fused --launch-skip 16 --launch-count 1 -f -o gpurun_out/${TAG}_ncu python bench.py --steps 1 --warmup 0 --no-cpu > gpurun_out/ncu_${TAG}.log 2>&1
  ls -la gpurun_out/${TAG}_ncu.ncu-rep
fi
