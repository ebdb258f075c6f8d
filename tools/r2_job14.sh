#!/bin/bash
# 1/8 shard of the 4K stress frame on one GPU: per-kernel times, and the frame with 1 / 2 / 3 launch chains
mkdir -p gpurun_out/j14
P="python tools/perf_probe.py --w 3840 --h 2160 --world 8 --rank 0 --reps 8"
$P > gpurun_out/j14/shard_profile.json 2>&1
for k in 1 2 3; do $P --nop --no-instrument --subframes $k > gpurun_out/j14/shard_nop_k$k.json 2>&1; done
for f in gpurun_out/j14/*.json; do echo $f; cut -c1-900 $f; done
