#!/bin/bash
run() { timeout 200 python tools/perf_probe.py --w 3840 --h 2160 --world $1 --rank 0 --nop --no-instrument --reps 5 \
      | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$2 world',$1,'ms',d['ms'])"; }
for w in 8 4 2 1; do
  for t in 0 1; do for s in 0 1; do
    RT_STEAL=$t RT_STEAL_SHADOW=$s run $w trace${t}_shadow${s}
  done; done
done
