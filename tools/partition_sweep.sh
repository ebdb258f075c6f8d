#!/bin/bash
run() { timeout 200 python tools/perf_probe.py --w 3840 --h 2160 --world $1 --rank 0 --nop --no-instrument --reps 5 \
      | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$2 world',$1,'ms',d['ms'])"; }
for w in 8 1; do
  run $w base
  for tm in 3 4 5; do for sm in 2 3 4; do
    RT_GRID_MULT=$tm RT_SHADOW_GRID_MULT=$sm run $w trace${tm}_shadow${sm}
  done; done
done
