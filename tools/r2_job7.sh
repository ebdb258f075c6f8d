#!/bin/bash
# Round 2, job 7: the octant-ordered 8-wide closest-hit walk: parity suite, then A/B against the 4-wide walk.
python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q > gpurun_out/r2_j7_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_j7_pytest.log
tail -4 gpurun_out/r2_j7_pytest.log
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 4"
for v in A W4 W8B6 A; do
  if [ $v = A ]; then unset RT_B200_LIB; else export RT_B200_LIB=simple_raytracer_b200/lib/variants/$v.so; fi
  echo "== $v profile" >> gpurun_out/r2_j7_ab.txt; $P >> gpurun_out/r2_j7_ab.txt 2>> gpurun_out/err_j7.log
  echo "== $v overlapped" >> gpurun_out/r2_j7_ab.txt; $P --nop --no-instrument >> gpurun_out/r2_j7_ab.txt 2>> gpurun_out/err_j7.log
  echo "== $v shard8" >> gpurun_out/r2_j7_ab.txt; $P --nop --no-instrument --world 8 --rank 3 >> gpurun_out/r2_j7_ab.txt 2>> gpurun_out/err_j7.log
done
