#!/bin/bash
# quick iteration: all GPU op/model tests, then cfg2 bench unfused + fused. usage: gpu_iter3.sh TAG
TAG=${1:-it}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${TAG}_pytest.log
IRB_FUSED_FFN=0 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/${TAG}_bench_unfused.json 2> gpurun_out/${TAG}_bench_unfused.err; echo "bench0 rc=$?"
IRB_FUSED_FFN=1 timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu --top 40 > gpurun_out/${TAG}_bench_fused.json 2> gpurun_out/${TAG}_bench_fused.err; echo "bench1 rc=$?"
python tools/show_bench.py gpurun_out/${TAG}_bench_unfused.json 2>/dev/null | head -22
python tools/show_bench.py gpurun_out/${TAG}_bench_fused.json 2>/dev/null | head -8
