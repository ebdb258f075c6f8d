#!/bin/bash
# inner / leaf phase thresholds of the persistent walks, shard and whole frame.
for shard in "--world 8 --rank 3" ""; do
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 7 --no-instrument --nop $shard"
echo "---- ${shard:-whole frame}"
run() { echo -n "$* : "; env "$@" timeout 200 $P | python -c "import sys,json; print(json.loads(sys.stdin.read().strip().splitlines()[-1])['ms'])"; }
run A=0
for lm in 2 4 6 8 12; do run RT_LEAF_MIN_LANES=$lm; done
for im in 6 8 12; do run RT_INNER_MIN_LANES=$im; done
for lm in 3 4 6; do for im in 6 8; do run RT_LEAF_MIN_LANES=$lm RT_INNER_MIN_LANES=$im; done; done
done
