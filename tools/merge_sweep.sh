#!/bin/bash
# Late generations' shadow chains in one merged any-hit launch (RT_SHADOW_MERGE_FROM).
for m in 0 2; do echo "parity, merge from $m:"; RT_SHADOW_MERGE_FROM=$m timeout 300 python tools/parity_quick.py; done
for shard in "" "--world 2 --rank 1" "--world 4 --rank 1" "--world 8 --rank 3"; do
  P="python tools/perf_probe.py --w 3840 --h 2160 --reps 7 --no-instrument --nop $shard"
  echo "---- ${shard:-whole frame}"
  for m in 99 0 1 2 3; do
    echo -n "merge from $m: "; RT_SHADOW_MERGE_FROM=$m timeout 200 $P | python -c "import sys,json; print(json.loads(sys.stdin.read().strip().splitlines()[-1])['ms'])"
  done
done
