#!/bin/bash
brief() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ms',d['ms'],'trace',d['trace_ms'][:1],'shadow',d['shadow_ms'][:1],'shade',d['shade_ms'][:1],'launches',d['launches'])"; }
for shard in "" "--world 8 --rank 3"; do
  P="python tools/perf_probe.py --w 3840 --h 2160 --reps 5 --no-instrument $shard"
  echo "---- ${shard:-whole frame}"
  for m in 7 3; do for se in 12 16 24; do
  echo "== profiled (serial) fused frame, mode $m, every $se"; RT_FUSED=1 RT_HINT_GROUP=32 RT_HINT_MODE=$m RT_SPAWN_EVERY=$se timeout 200 $P | brief
  done; done
done
