#!/bin/bash
# Round 2, job 5: graph replay test + suite subset, small-scene benches with graphs, level_kernel occupancy A/B.
python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r2_j5_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_j5_pytest.log
for w in four_spheres_512 test7_1080p test7_4k house_4k; do
  python bench.py --workload $w --steps 20 --no-e2e --no-cpu > gpurun_out/r2_j5_$w.json 2>> gpurun_out/err_j5.log
done
for v in level6 level8; do
  for w in four_spheres_512 test7_4k; do
    RT_B200_LIB=simple_raytracer_b200/lib/variants/$v.so python bench.py --workload $w --steps 20 --no-e2e --no-cpu > gpurun_out/r2_j5_${v}_$w.json 2>> gpurun_out/err_j5.log
  done
done
python tools/perf_probe.py --w 3840 --h 2160 --reps 6 --nop --no-instrument > gpurun_out/r2_j5_probe_4k.json 2>> gpurun_out/err_j5.log
python tools/perf_probe.py --w 3840 --h 2160 --reps 6 --nop --no-instrument --world 8 --rank 3 > gpurun_out/r2_j5_probe_shard8.json 2>> gpurun_out/err_j5.log
tail -3 gpurun_out/r2_j5_pytest.log
