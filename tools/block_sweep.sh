#!/bin/bash
# Block size of the walk kernels (128 / 64 / 32 threads): overlapped frame, whole and as shards.
V=$PWD/simple_raytracer_b200/lib/variants
for shard in "" "--world 2 --rank 1" "--world 4 --rank 1" "--world 8 --rank 3"; do
  P="python tools/perf_probe.py --w 3840 --h 2160 --reps 7 --no-instrument --nop $shard"
  echo "---- ${shard:-whole frame}"
  for lib in "" $V/librt_b200_wb64.so $V/librt_b200_wb32.so; do
    echo -n "${lib##*/} : "; RT_B200_LIB=$lib timeout 200 $P | python -c "import sys,json; print(json.loads(sys.stdin.read().strip().splitlines()[-1])['ms'])"
  done
done
