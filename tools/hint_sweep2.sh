#!/bin/bash
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 4"
brief() { python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('ms',d['ms'],'shadow_ms',d['shadow_ms'],'trace_ms',d['trace_ms'],'walked',d.get('shadow_walked'),'hits',d.get('hint_hits'),'boxes/ray',d.get('boxes_per_ray'),'prims/ray',d.get('prims_per_ray'))"; }
run() { echo "== $*"; env "$@" timeout 300 $P | brief; }
run RT_HINT=0
run RT_HINT=1 RT_HINT_MODE=0
run RT_HINT=1 RT_HINT_MODE=0 RT_HINT_GROUP=32
run RT_HINT=1 RT_HINT_KEEP=1
run RT_HINT=1 RT_HINT_KEEP=1 RT_HINT_PASSES=1
run RT_HINT=1 RT_HINT_PASSES=1
run RT_HINT=1 RT_HINT_GROUP=32
run RT_HINT=1 RT_HINT_GROUP=32 RT_HINT_PASSES=64 RT_HINT_MIN_CHUNK_K=128
run RT_HINT=1 RT_HINT_GROUP=256
run RT_HINT=1 RT_HINT_GROUP=3840
run RT_HINT=1 RT_HINT_GROUP=32 RT_REFILL_LANES=2
run RT_HINT=1 RT_HINT_GROUP=32 RT_REFILL_LANES=12
run RT_HINT=1 RT_HINT_GROUP=32 RT_INNER_MIN_LANES=8
run RT_HINT=1 RT_HINT_GROUP=32 RT_INNER_MIN_LANES=1
