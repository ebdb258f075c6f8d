#!/bin/bash
# after the occluder maps: smoke, the whole GPU suite, every workload's bench line, the reference arm, then the
# ncu launch list and one full capture of the default bench (each only after the same command exited 0 without ncu)
mkdir -p gpurun_out/j19
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/j19/smoke.log 2>&1; echo "smoke rc $?"
python -m pytest tests -m gpu -x -q > gpurun_out/j19/pytest.log 2>&1; echo "pytest rc $?"; tail -4 gpurun_out/j19/pytest.log
for w in stress_1M_4k stress_1M_8k mesh_earth_1080p test7_1080p test7_4k house_4k four_spheres_512; do
  python bench.py --workload $w --steps 20 --warmup 5 > gpurun_out/j19/r2_bench_1gpu_$w.json 2> gpurun_out/j19/r2_bench_1gpu_$w.err
  echo "$w rc=$?"; cut -c1-260 gpurun_out/j19/r2_bench_1gpu_$w.json
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/j19/r2_bench_reference_arm.json 2> gpurun_out/j19/ref.err; echo "ref rc=$?"
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
$B > gpurun_out/j19/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/j19/r2_bench_launches.csv $B > gpurun_out/j19/ncu_launches.log 2>&1
$B > gpurun_out/j19/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -f -k 'regex:bvh_(trace|shadow|shadow_settle|colour|spawn)_kernel' -s 70 -c 14 -o gpurun_out/j19/r2_bvh_full $B > gpurun_out/j19/ncu_full.log 2>&1
python tools/exactness_report.py gpurun_out/j19/r2_exactness.txt > /dev/null 2>&1
ls -la gpurun_out/j19 | tail -30
