#!/bin/bash
# Round-end measurements on one B200: tests, bench lines, ncu launch list and full capture.
set -x
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py > gpurun_out/r1_bench_1gpu.json 2> gpurun_out/bench.err || tail -5 gpurun_out/bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r1_bench_reference_arm.json 2>> gpurun_out/bench.err
for w in stress_1M_8k mesh_earth_1080p test7_4k four_spheres_512; do
  python bench.py --workload $w --no-cpu --steps 10 > gpurun_out/r1_bench_1gpu_$w.json 2>> gpurun_out/bench.err
done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_bench_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -f -k 'regex:bvh_(trace|shadow|shade)_kernel' -s 40 -c 8 -o gpurun_out/r1_bvh_full_d \
    python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -20
