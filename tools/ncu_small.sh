#!/bin/bash
# Launch list + full capture of the brute-force (small scene) path: test7 at 4K.
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/test7_launches.csv \
    python bench.py --workload test7_4k --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_s1.log 2>&1
ncu --set full --clock-control none --import-source on -f -k 'regex:level_kernel|combine_kernel' -s 33 -c 6 -o gpurun_out/test7_full \
    python bench.py --workload test7_4k --steps 2 --warmup 3 --no-e2e --no-cpu > gpurun_out/ncu_s2.log 2>&1
tail -2 gpurun_out/ncu_s2.log
