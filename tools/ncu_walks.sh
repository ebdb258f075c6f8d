#!/bin/bash
# Full ncu captures (source counters) of single walk launches of the 4K stress frame.
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 1 --no-instrument"
N="ncu --set full --import-source on --clock-control none -f"
RT_HINT=0 $N -k regex:bvh_shadow_kernel -s 5 -c 1 -o gpurun_out/shadow_g0_nohint $P > gpurun_out/ncu_a.log 2>&1
RT_HINT=1 RT_HINT_GROUP=32 $N -k regex:bvh_shadow_kernel -s 5 -c 1 -o gpurun_out/shadow_g0_hint $P > gpurun_out/ncu_b.log 2>&1
RT_HINT=0 $N -k regex:bvh_trace_kernel -s 6 -c 1 -o gpurun_out/trace_g1 $P > gpurun_out/ncu_c.log 2>&1
RT_HINT=0 $N -k regex:bvh_trace_kernel -s 5 -c 1 -o gpurun_out/trace_g0 $P > gpurun_out/ncu_d.log 2>&1
ls -la gpurun_out/
