#!/bin/bash
# Fused-GDFN bring-up: op parity, model parity, bench A/B (IRB_FUSED_FFN=0 vs 1). usage: gpu_ffn.sh TAG
TAG=${1:-ffn}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k "gdfn_fused" > gpurun_out/${TAG}_op.log 2>&1; echo "op rc=$?"; tail -15 gpurun_out/${TAG}_op.log
timeout 900 python -m pytest tests/test_gpu_models.py tests/test_gpu_tools.py -m gpu -x -q > gpurun_out/${TAG}_models.log 2>&1; echo "models rc=$?"; tail -8 gpurun_out/${TAG}_models.log
IRB_FUSED_FFN=0 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/${TAG}_bench_unfused.json 2> gpurun_out/${TAG}_bench_unfused.err; echo "bench0 rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/${TAG}_bench_fused.json 2> gpurun_out/${TAG}_bench_fused.err; echo "bench1 rc=$?"
python tools/show_bench.py gpurun_out/${TAG}_bench_unfused.json | head -12
python tools/show_bench.py gpurun_out/${TAG}_bench_fused.json | head -30
