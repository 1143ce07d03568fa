#!/bin/bash
# The one GPU-box driver script (run it through gpurun). usage: tools/gpu_round.sh TAG [steps...]
#   steps: test smoke bench[:wl1,wl2,...] ref launches ncu:<kernel-regex>:<skip>[:<workload>] ab:<lib_a>:<lib_b>[:<wl>]
#          graph_ab[:wl,...]
# Everything is written under gpurun_out/ with the TAG prefix; copy what should be judged into profiles/.
TAG=${1:-rXX}; shift
STEPS=${@:-test smoke bench ref}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/${TAG}_gpu.txt 2>&1; nproc >> gpurun_out/${TAG}_gpu.txt
for st in $STEPS; do
  IFS=: read kind a b c <<< "$st"
  case $kind in
    test)
      timeout 1500 python -m pytest tests -m gpu -x -q -rP > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log
      grep -h "^\[parity\]" gpurun_out/${TAG}_pytest_gpu.log | sort -u > gpurun_out/${TAG}_parity.txt
      tail -3 gpurun_out/${TAG}_pytest_gpu.log;;
    smoke)
      timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${TAG}_smoke.log; tail -2 gpurun_out/${TAG}_smoke.log;;
    bench)
      for wl in $(echo ${a:-cfg2,cfg5,swinir_dn,cfg3_1img,cfg1} | tr , ' '); do
        extra="--no-cpu"; [ $wl = cfg2 ] && extra=""
        timeout 900 python bench.py --workload $wl --warmup 3 --top 60 $extra > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err; echo "bench $wl rc=$?"
      done;;
    benchcpu)   # with the parity + cpu_baseline legs
      for wl in $(echo ${a:-cfg2} | tr , ' '); do
        timeout 1500 python bench.py --workload $wl --warmup 3 --top 60 > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err; echo "bench $wl rc=$?"
      done;;
    optab)   # optab:<key>:<workloads>  library option 0 vs 1 on one box, interleaved twice
      for rep in 1 2; do for v in 0 1; do for wl in $(echo ${b:-cfg2} | tr , ' '); do
        timeout 600 python bench.py --workload $wl --warmup 3 --no-cpu --opt $a=$v --top 40 > gpurun_out/${TAG}_opt_${a}${v}_${wl}_$rep.json 2>> gpurun_out/${TAG}_optab.err; echo "opt $a=$v $wl rc=$?"
      done; done; done;;
    pytest)  # pytest:<k-expression>
      timeout 1200 python -m pytest tests -m gpu -x -q -rP -k "$a" > gpurun_out/${TAG}_pytest_$a.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_$a.log; tail -15 gpurun_out/${TAG}_pytest_$a.log;;
    graph_ab)
      for wl in $(echo ${a:-cfg2,swinir_dn} | tr , ' '); do for g in 0 1; do
        timeout 600 python bench.py --workload $wl --warmup 3 --no-cpu --graph $g --top 5 > gpurun_out/${TAG}_graph${g}_$wl.json 2> gpurun_out/${TAG}_graph${g}_$wl.err; echo "graph$g $wl rc=$?"
      done; done;;
    ref)
      timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_reference_arm.json 2>&1;;
    launches)
      CMD="python bench.py --workload ${a:-cfg2} --steps 1 --warmup 3 --no-cpu --graph 0"
      timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/${TAG}_launches_${a:-cfg2}.csv $CMD > gpurun_out/ncu.log 2>&1;;
    ncu)   # ncu:<kernel regex>:<launch-skip>[:workload]
      CMD="python bench.py --workload ${c:-cfg2} --steps 1 --warmup 0 --no-cpu --graph 0"
      timeout 600 $CMD > gpurun_out/plain2.log 2>&1 &&
      timeout 900 ncu --set full --import-source on --clock-control none -k regex:$a --launch-skip ${b:-0} --launch-count 1 -f -o gpurun_out/${TAG}_ncu_${a}_${b:-0} $CMD > gpurun_out/ncu_full.log 2>&1
      ncu -i gpurun_out/${TAG}_ncu_${a}_${b:-0}.ncu-rep --page details > gpurun_out/${TAG}_ncu_full_${a}_${b:-0}.txt 2>&1;;
    ab)    # ab:<lib a>:<lib b>[:workload]  two builds of the library on one box, interleaved
      for rep in 1 2; do for lib in $a $b; do
        IRB_LIB=$lib timeout 600 python bench.py --workload ${c:-cfg2} --warmup 3 --no-cpu --top 40 > gpurun_out/${TAG}_ab_$(basename $lib .so)_${c:-cfg2}_$rep.json 2>> gpurun_out/${TAG}_ab.err
      done; done;;
    *) echo "unknown step $st";;
  esac
done
echo done
