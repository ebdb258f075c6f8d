#!/bin/bash
# occluder maps: parity, per-kernel times, resolution sweep (lab build)
mkdir -p gpurun_out/j18
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_multigpu_abi.py -x -q > gpurun_out/j18/pytest.log 2>&1
echo "pytest rc $?"; tail -5 gpurun_out/j18/pytest.log
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 6"
timeout 300 $P > gpurun_out/j18/full_profile.json 2>&1
timeout 300 $P --nop --no-instrument > gpurun_out/j18/full_nop.json 2>&1
timeout 300 $P --world 8 --rank 0 --nop --no-instrument > gpurun_out/j18/shard_nop.json 2>&1
for f in gpurun_out/j18/*.json; do echo $f; cut -c1-700 $f; done
export RT_B200_LIB=$PWD/simple_raytracer_b200/lib/librt_b200_lab.so
for res in 1024 2048 4096 8192; do
  echo "== RT_SMAP_MAX_RES=$res"
  RT_SMAP_PER_SQRT=16 RT_SMAP_MAX_RES=$res timeout 300 $P --reps 4 2>&1 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print({k:d[k] for k in ('ms','shadow_ms','shadow_walked','first_render_s')})"
  RT_SMAP_PER_SQRT=16 RT_SMAP_MAX_RES=$res timeout 300 $P --reps 6 --nop --no-instrument 2>&1 | cut -c1-80
done
