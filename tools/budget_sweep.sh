#!/bin/bash
run() { timeout 200 python tools/perf_probe.py --w 3840 --h 2160 --world $1 --rank 0 --nop --no-instrument --reps 5 \
      | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$2 world',$1,'ms',d['ms'])"; }
for w in 8 4 1; do
  for b in 0 32 64 128 256 512; do
    RT_SHADOW_JOB_BUDGET=$b run $w budget$b
  done
done
