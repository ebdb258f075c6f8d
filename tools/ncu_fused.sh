#!/bin/bash
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 1 --no-instrument"
N="ncu --set full --import-source on --clock-control none -f"
RT_FUSED=1 RT_SPAWN_EVERY=16 timeout 600 $N -k regex:bvh_trace_all_kernel -s 1 -c 1 -o gpurun_out/trace_all $P > gpurun_out/ncu_e.log 2>&1
tail -3 gpurun_out/ncu_e.log
