set -x
P=/root/repo/image_restoration_agent_b200
for rep in 1 2; do for v in p0 ns0 ns1 ns2; do
  lib=$P/libirb200_$v.so
  for wl in swinir_dn hat_x4; do
    IRB_LIB=$lib timeout 600 python bench.py --workload $wl --warmup 3 --no-cpu --top 12 > gpurun_out/r02_s71_${v}_${wl}_$rep.json 2>> gpurun_out/r02_s71.err
  done
done; done
echo done
