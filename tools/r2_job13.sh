#!/bin/bash
# refresh the bench lines of the other workloads with the final kernels (spawn kernel timed on its own)
mkdir -p gpurun_out/j13
for w in stress_1M_8k mesh_earth_1080p test7_1080p test7_4k house_4k four_spheres_512; do
  python bench.py --workload $w --steps 20 --warmup 5 > gpurun_out/j13/r2_bench_1gpu_$w.json 2> gpurun_out/j13/r2_bench_1gpu_$w.err
  echo "$w rc=$?"; cut -c1-300 gpurun_out/j13/r2_bench_1gpu_$w.json
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/j13/r2_bench_reference_arm.json 2> gpurun_out/j13/ref.err; echo "ref rc=$?"
