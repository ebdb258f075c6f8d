#!/bin/bash
# Occluder hints of the any-hit kernel: off / on, pass and group sizes, 7 vs 8 blocks per SM.
# 4K stress frame, whole and as a 1/8 shard; --nop = the overlapped frame as bench.py times it.
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 5"
run() { echo "== $*"; env "$@" timeout 200 $P --nop --no-instrument | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms'])"; }
echo "-- whole frame, overlapped"
run RT_HINT=0
run RT_HINT=1
run RT_HINT=1 RT_HINT_GROUP=4
run RT_HINT=1 RT_HINT_GROUP=32
run RT_HINT=1 RT_HINT_PASSES=8
run RT_HINT=1 RT_HINT_PASSES=32 RT_HINT_MIN_CHUNK_K=192
run RT_HINT=1 RT_HINT_PASSES=1
run RT_HINT=1 RT_B200_LIB=$PWD/simple_raytracer_b200/lib/variants/librt_b200_sh7.so
echo "-- whole frame, per kernel + counters"
RT_HINT=0 timeout 300 $P
RT_HINT=1 timeout 300 $P
RT_HINT=1 RT_HINT_GROUP=32 timeout 300 $P
echo "-- 1/8 shard, overlapped"
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 5 --world 8 --rank 3"
run RT_HINT=0
run RT_HINT=1
run RT_HINT=1 RT_HINT_MIN_CHUNK_K=128
run RT_HINT=1 RT_HINT_MIN_CHUNK_K=64
run RT_HINT=1 RT_HINT_GROUP=32 RT_HINT_MIN_CHUNK_K=128
