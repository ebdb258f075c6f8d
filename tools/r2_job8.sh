#!/bin/bash
# Round 2, job 8: prefetch of the second-nearest child's node in the 4-wide step (P1: into L2, P2: into L1).
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 4 --no-instrument"
for v in A P1 P2 A; do
  if [ $v = A ]; then unset RT_B200_LIB; else export RT_B200_LIB=simple_raytracer_b200/lib/variants/$v.so; fi
  echo "== $v profile" >> gpurun_out/r2_j8_ab.txt; $P >> gpurun_out/r2_j8_ab.txt 2>> gpurun_out/err_j8.log
  echo "== $v overlapped" >> gpurun_out/r2_j8_ab.txt; $P --nop >> gpurun_out/r2_j8_ab.txt 2>> gpurun_out/err_j8.log
  echo "== $v shard8" >> gpurun_out/r2_j8_ab.txt; $P --nop --world 8 --rank 3 >> gpurun_out/r2_j8_ab.txt 2>> gpurun_out/err_j8.log
done
