#!/bin/bash
run() { timeout 200 python tools/perf_probe.py --w 3840 --h 2160 --world $1 --rank 0 --nop --no-instrument --subframes $2 --reps 5 \
      | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$3 world',$1,'subframes',$2,'ms',d['ms'])"; }
for w in 8 1; do
  for k in 1 2 4; do
    CUDA_DEVICE_MAX_CONNECTIONS=32 run $w $k conn32
  done
  for m in 3 5 7 8 12; do
    RT_GRID_MULT=$m run $w 1 gridmult$m
    RT_GRID_MULT=$m run $w 2 gridmult$m
  done
done
