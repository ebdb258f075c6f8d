#!/bin/bash
# Round 2, job 6: occupancy A/B of the walk kernels (registers against resident warps).
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 4 --no-instrument"
for v in A T8 T9 S10 S12 A; do
  if [ $v = A ]; then unset RT_B200_LIB; else export RT_B200_LIB=simple_raytracer_b200/lib/variants/$v.so; fi
  echo "== $v profile" >> gpurun_out/r2_j6_ab.txt; $P >> gpurun_out/r2_j6_ab.txt 2>> gpurun_out/err_j6.log
  echo "== $v overlapped" >> gpurun_out/r2_j6_ab.txt; $P --nop >> gpurun_out/r2_j6_ab.txt 2>> gpurun_out/err_j6.log
  echo "== $v shard8" >> gpurun_out/r2_j6_ab.txt; $P --nop --world 8 --rank 3 >> gpurun_out/r2_j6_ab.txt 2>> gpurun_out/err_j6.log
done
