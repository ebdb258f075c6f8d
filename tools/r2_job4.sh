#!/bin/bash
# Round 2, job 4: suite + exactness after the screening fast paths; small-scene benches; ncu of level_kernel.
python -m pytest tests -m gpu -x -q > gpurun_out/r2_j4_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_j4_pytest.log
python tools/exactness_report.py gpurun_out/r2_j4_exactness.txt > /dev/null 2> gpurun_out/err_ex.log
B="python bench.py --workload test7_4k --steps 10 --warmup 3 --no-e2e --no-cpu"
$B > gpurun_out/r2_j4_test7_4k.json 2> gpurun_out/err1.log &&
ncu --set full --clock-control none --import-source on -f -k regex:level_kernel -s 15 -c 5 -o gpurun_out/r2_level_test7_b $B > gpurun_out/ncu4.log 2>&1
python bench.py --workload four_spheres_512 --steps 20 --no-e2e --no-cpu > gpurun_out/r2_j4_four_spheres.json 2> gpurun_out/err2.log
python bench.py --workload house_4k --steps 10 --no-e2e --no-cpu > gpurun_out/r2_j4_house_4k.json 2> gpurun_out/err3.log
python bench.py --steps 10 --no-cpu > gpurun_out/r2_j4_stress_4k.json 2> gpurun_out/err5.log
tail -3 gpurun_out/r2_j4_pytest.log; tail -1 gpurun_out/r2_j4_exactness.txt
