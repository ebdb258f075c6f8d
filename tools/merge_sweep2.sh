#!/bin/bash
# Two-phase frame (trace chains, then merged any-hit launches) with K interleaved sub-frame chains.
for shard in "--world 8 --rank 3" "--world 4 --rank 1" "--world 2 --rank 1" ""; do
  P="python tools/perf_probe.py --w 3840 --h 2160 --reps 7 --no-instrument --nop $shard"
  echo "---- ${shard:-whole frame}"
  for k in 1 2 3 4; do for m in 0 99; do
    echo -n "subframes $k merge from $m: "; RT_SUBFRAMES=$k RT_SHADOW_MERGE_FROM=$m timeout 200 $P | python -c "import sys,json; print(json.loads(sys.stdin.read().strip().splitlines()[-1])['ms'])"
  done; done
done
