#!/bin/bash
# Round checkpoint: parity tests, smoke, bench on all workloads (+ reference arm), ncu launch list and one ncu --set full
# capture of the dominant kernel of the default bench. usage: gpu_round.sh TAG
set -x
TAG=${1:-rXX}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${TAG}_gpu.txt 2>&1; nproc >> gpurun_out/${TAG}_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -3 gpurun_out/${TAG}_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${TAG}_smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 --top 60 > gpurun_out/${TAG}_bench_cfg2.json 2> gpurun_out/${TAG}_bench_cfg2.err; echo "bench rc=$?"
for wl in cfg5 cfg3 cfg1 swinir_dn; do
  timeout 600 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu --top 60 > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err; echo "bench $wl rc=$?"
done
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference_arm.json 2>&1
CMD="python bench.py --steps 1 --warmup 3 --no-cpu"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${TAG}_launches_cfg2.csv $CMD > gpurun_out/ncu.log 2>&1
# dominant kernel (dw3x3 gate, decoder level 1 = launch 36.. of 44) with the full set
timeout 900 ncu --set full --import-source on --clock-control none -k regex:dw3x3_tma --launch-skip 73 --launch-count 1 -f -o gpurun_out/${TAG}_dw_gate python bench.py --steps 1 --warmup 0 --no-cpu > gpurun_out/ncu_full.log 2>&1
echo done
