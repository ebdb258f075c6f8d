#!/bin/bash
# The current any-hit kernel against the build before occluder hints (variants/librt_b200_old.so).
V=$PWD/simple_raytracer_b200/lib/variants
brief() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ms',d['ms'],'shadow_ms',d['shadow_ms'],sum(d['shadow_ms']),'trace',round(sum(d['trace_ms']),2),'shade',round(sum(d['shade_ms']),2))"; }
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 5 --no-instrument"
echo "old, profiled"; RT_B200_LIB=$V/librt_b200_old.so timeout 200 $P | brief
echo "new hints off, profiled"; RT_HINT=0 timeout 200 $P | brief
echo "new hints on, profiled"; timeout 200 $P | brief
for shard in "" "--world 8 --rank 3"; do
echo "--- overlapped ${shard:-whole}"
echo -n "old: "; RT_B200_LIB=$V/librt_b200_old.so timeout 200 $P --nop $shard | python -c "import sys,json; print(json.loads(sys.stdin.read().strip().splitlines()[-1])['ms'])"
echo -n "new hints off: "; RT_HINT=0 timeout 200 $P --nop $shard | python -c "import sys,json; print(json.loads(sys.stdin.read().strip().splitlines()[-1])['ms'])"
echo -n "new hints on: "; timeout 200 $P --nop $shard | python -c "import sys,json; print(json.loads(sys.stdin.read().strip().splitlines()[-1])['ms'])"
done
