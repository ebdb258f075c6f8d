#!/bin/bash
# sub-frame chains: frame time of the 4K stress frame, whole and as a 1/2, 1/4, 1/8 shard
for w in 8 4 2 1; do
  for k in 1 2 3 4 6 8; do
    timeout 200 python tools/perf_probe.py --w 3840 --h 2160 --world $w --rank 0 --nop --no-instrument --subframes $k --reps 5 \
      | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('world',$w,'subframes',$k,'ms',d['ms'],'launches',d['launches'])"
  done
done
