#!/bin/bash
# one full ncu capture covering a whole frame of the default bench (both chains, all generations); the report is
# too large to travel (> 64 MiB): its raw page is exported on the box
mkdir -p gpurun_out/j23
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu"
$B > gpurun_out/j23/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -f -k 'regex:bvh_(trace|shadow|shadow_settle|shadow_probe|colour|spawn)_kernel' -s 120 -c 44 -o /tmp/r2_bvh_full $B > gpurun_out/j23/ncu_full.log 2>&1
ncu -i /tmp/r2_bvh_full.ncu-rep --page raw --csv > gpurun_out/j23/r2_bvh_full_raw.csv 2> gpurun_out/j23/export.err
ls -la gpurun_out/j23; tail -3 gpurun_out/j23/ncu_full.log
