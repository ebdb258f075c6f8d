#!/bin/bash
# Round-end measurements on ONE B200: every workload's bench line, the reference arm, the ncu launch list and
# one full capture of the default bench (each ncu run only after the same command exited 0 without ncu).
mkdir -p gpurun_out
for w in stress_1M_4k stress_1M_8k mesh_earth_1080p test7_1080p test7_4k house_4k four_spheres_512; do
  python bench.py --workload $w --steps 20 --warmup 5 > gpurun_out/r2_bench_1gpu_$w.json 2> gpurun_out/r2_bench_1gpu_$w.err
  echo "$w rc=$?"
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2> gpurun_out/r2_bench_reference_arm.err
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
$B > gpurun_out/r2_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_bench_launches.csv $B > gpurun_out/r2_ncu_launches.log 2>&1
$B > gpurun_out/r2_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -f -k 'regex:bvh_(trace|shadow|colour|spawn)_kernel' -s 60 -c 12 -o gpurun_out/r2_bvh_full $B > gpurun_out/r2_ncu_full.log 2>&1
python tools/exactness_report.py gpurun_out/r2_exactness.txt > /dev/null 2>&1
ls -la gpurun_out | tail -30
