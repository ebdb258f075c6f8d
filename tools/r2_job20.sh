#!/bin/bash
# cube maps for point lights: parity, then frames with point lights first / second
mkdir -p gpurun_out/j20
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_multigpu_abi.py tests/test_gpu_fullsize.py -x -q > gpurun_out/j20/pytest.log 2>&1
echo "pytest rc $?"; tail -5 gpurun_out/j20/pytest.log
python - <<'PY'
import sys, time, json
sys.path.insert(0, '.')
import numpy as np
from simple_raytracer_b200 import abi, scenegen
from simple_raytracer_b200.renderer import DeviceScene
def run(name, sc):
    with DeviceScene(sc) as ds:
        t0 = time.time(); ds.render(accel=abi.RT_ACCEL_LBVH); first = time.time() - t0
        st = ds.render(accel=abi.RT_ACCEL_LBVH, instrument=True)
        best = min(ds.render(accel=abi.RT_ACCEL_LBVH).render_ms for _ in range(8))
        print(json.dumps({"scene": name, "ms": round(best, 3), "first_render_s": round(first, 3), "shadow_ref": int(st.rays_shadow),
                          "walked": int(st.shadow_rays_walked), "map_hits": int(st.shadow_map_hits)}), flush=True)
sc = scenegen.stress_scene(1_000_000, 1000, 3840, 2160)
run("stress 4K", sc)
s2 = sc.copy(); s2.lights = s2.lights[::-1].copy(); s2._invalidate()
run("stress 4K, point lights first", s2)
run("mesh_earth_1080p", scenegen.mesh_earth(512, 256, 1920, 1080))
PY
