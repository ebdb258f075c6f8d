#!/bin/bash
# L1 policy experiments of the walk kernels: primitive records loaded no_allocate / evict_first,
# 16 / 32 shared-memory stack entries, L1 / shared-memory carve-out.  4K stress frame.
V=$PWD/simple_raytracer_b200/lib/variants
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 5 --no-instrument"
brief() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ms',d['ms'],'shadow',round(sum(d['shadow_ms']),2),'trace',round(sum(d['trace_ms']),2),'shade',round(sum(d['shade_ms']),2))"; }
ov() { python -c "import sys,json; print('overlapped', json.loads(sys.stdin.read().strip().splitlines()[-1])['ms'])"; }
for lib in "" $V/librt_b200_pl1.so $V/librt_b200_pl2.so $V/librt_b200_ss16.so $V/librt_b200_ss32.so; do
  echo "== ${lib##*/}"; RT_B200_LIB=$lib timeout 200 $P | brief; RT_B200_LIB=$lib timeout 200 $P --nop | ov
done
for c in 25 40 50 75 100; do
  echo "== carveout $c"; RT_CARVEOUT=$c timeout 200 $P | brief; RT_CARVEOUT=$c timeout 200 $P --nop | ov
done
