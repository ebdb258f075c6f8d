#!/bin/bash
# Remaining scheduling knobs re-checked in the two-phase frame with the new thresholds.
for shard in "" "--world 8 --rank 3"; do
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 7 --no-instrument --nop $shard"
echo "---- ${shard:-whole frame}"
run() { echo -n "$* : "; env "$@" timeout 200 $P | python -c "import sys,json; print(json.loads(sys.stdin.read().strip().splitlines()[-1])['ms'])"; }
run A=0
run RT_GRID_MULT=8
run RT_GRID_MULT=12
run RT_GRID_MULT=32
run RT_STEAL_MAX_KRAYS=0
run RT_STEAL_MAX_KRAYS=5000
run RT_STEAL_MAX_KRAYS=20000
run RT_STEAL_SHADOW=0
run RT_STEAL_SHADOW=1
run RT_SUBFRAMES=1
run RT_SUBFRAMES=2
run RT_SUBFRAMES=3
run RT_SHADOW_REFILL_LANES=3
run RT_SHADOW_REFILL_LANES=10
run RT_BVH8=0
done
