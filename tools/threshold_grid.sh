#!/bin/bash
# Profiled (serial) frames: each kernel's time depends only on its own thresholds, so one grid over
# (inner_min, leaf_min, refill) tunes the closest-hit and the any-hit kernel at once.
for shard in "" "--world 8 --rank 3"; do
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 5 --no-instrument $shard"
echo "---- ${shard:-whole frame}   inner leaf refill : trace_ms shadow_ms"
for im in 4 6 8 10; do for lm in 1 4 6 8 12; do for rf in 6; do
  echo -n "$im $lm $rf : "; RT_INNER_MIN_LANES=$im RT_LEAF_MIN_LANES=$lm RT_REFILL_LANES=$rf timeout 200 $P | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(sum(d['trace_ms']),3), round(sum(d['shadow_ms']),3))"
done; done; done
for rf in 2 4 8 12; do
  echo -n "8 6 $rf : "; RT_INNER_MIN_LANES=8 RT_LEAF_MIN_LANES=6 RT_REFILL_LANES=$rf timeout 200 $P | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(sum(d['trace_ms']),3), round(sum(d['shadow_ms']),3))"
done
done
