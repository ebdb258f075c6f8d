#!/bin/bash
# Round 2, job 3: A/B of the 4-wide step's sort (integer keys, branch-free push vs round-1 float network) and
# of the level-0 slot order (8x4 patches vs scanline), then ncu --set full of the walk kernels of build A.
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 4 --no-instrument"
for v in A B C D A B C D; do
  if [ $v = A ]; then unset RT_B200_LIB; else export RT_B200_LIB=simple_raytracer_b200/lib/variants/$v.so; fi
  echo "== $v profile" >> gpurun_out/r2_j3_ab.txt; $P >> gpurun_out/r2_j3_ab.txt 2>> gpurun_out/err_ab.log
  echo "== $v overlapped" >> gpurun_out/r2_j3_ab.txt; $P --nop >> gpurun_out/r2_j3_ab.txt 2>> gpurun_out/err_ab.log
  echo "== $v shard8" >> gpurun_out/r2_j3_ab.txt; $P --nop --world 8 --rank 3 >> gpurun_out/r2_j3_ab.txt 2>> gpurun_out/err_ab.log
done
unset RT_B200_LIB
Q="python tools/perf_probe.py --w 3840 --h 2160 --reps 1 --no-instrument --nop"
$Q > gpurun_out/r2_j3_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -f -k 'regex:bvh_(trace|shadow)_kernel' -s 6 -c 6 -o gpurun_out/r2_walks_a $Q > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/ncu3.log
