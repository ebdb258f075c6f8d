#!/bin/bash
# occluder maps of directional lights: parity subset, then the 4K stress frame and its 1/8 shard, per-kernel times
mkdir -p gpurun_out/j17
timeout 1200 python -m pytest tests/test_gpu_parity.py tests/test_multigpu_abi.py -x -q > gpurun_out/j17/pytest.log 2>&1
echo "pytest rc $?"; tail -5 gpurun_out/j17/pytest.log
P="python tools/perf_probe.py --w 3840 --h 2160 --reps 6"
timeout 300 $P > gpurun_out/j17/full_profile.json 2>&1
timeout 300 $P --nop --no-instrument > gpurun_out/j17/full_nop.json 2>&1
timeout 300 $P --world 8 --rank 0 --nop --no-instrument > gpurun_out/j17/shard_nop.json 2>&1
for f in gpurun_out/j17/*.json; do echo $f; cut -c1-1200 $f; done
python - <<'PY'
import sys, time
sys.path.insert(0, '.')
from simple_raytracer_b200 import abi, scenegen
from simple_raytracer_b200.renderer import DeviceScene
sc = scenegen.stress_scene(1_000_000, 1000, 3840, 2160)
t0 = time.time()
with DeviceScene(sc) as ds:
    st = ds.render(accel=abi.RT_ACCEL_LBVH, instrument=True)
    print("instrumented: shadow rays", st.rays_shadow, "walked", st.shadow_rays_walked, "map hits", st.shadow_map_hits, "nodes", st.nodes_total)
PY
