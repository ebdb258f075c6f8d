#!/bin/bash
# Walk-loop thresholds and work sharing at shard size (1/8 of the 4K stress frame), two-phase frame.
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 9 --no-instrument --nop --world 8 --rank 3"
run() { echo -n "$* : "; env "$@" timeout 200 $P | python -c "import sys,json; print(json.loads(sys.stdin.read().strip().splitlines()[-1])['ms'])"; }
run A=0
run RT_REFILL_LANES=2
run RT_REFILL_LANES=4
run RT_REFILL_LANES=10
run RT_INNER_MIN_LANES=2
run RT_INNER_MIN_LANES=8
run RT_LEAF_MIN_LANES=4
run RT_STEAL=0 RT_STEAL_SHADOW=0
run RT_STEAL=1 RT_STEAL_SHADOW=0
run RT_STEAL=0 RT_STEAL_SHADOW=1
run RT_GRID_MULT=7
run RT_GRID_MULT=4
run RT_SHADOW_GRID_MULT=8
run RT_SHADOW_GRID_MULT=4
run A=1
