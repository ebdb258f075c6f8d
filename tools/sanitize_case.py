#!/usr/bin/env python
"""A small frame through every kernel family (upload, LBVH build, both walk kernels with work
sharing, spawn / colour / combine, the brute-force level kernel, device PPM formatter) for
compute-sanitizer runs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simple_raytracer_b200 import abi, scenegen
from simple_raytracer_b200.renderer import DeviceScene

for sc in (scenegen.coverage_scene(48, 36), scenegen.stress_scene(3000, 40, 64, 48, tex_size=32)):
    with DeviceScene(sc) as ds:
        ds.render(accel=abi.RT_ACCEL_LBVH, work_sharing=1)
        a = ds.download()
        txt = ds.ppm_text()
        ds.render(accel=abi.RT_ACCEL_LBVH, work_sharing=2, subframes=3)
        ds.render(accel=abi.RT_ACCEL_LBVH, instrument=True)
        ds.render(accel=abi.RT_ACCEL_BRUTE)
        b = ds.download()
        assert (a[0] == b[0]).all() and (a[1] == b[1]).all() and len(txt) > 100
# graph capture + replay, the probes of both trees, N = 1 through the multi-GPU entry points
import numpy as np
from simple_raytracer_b200.renderer import MultiDeviceScene
sc = scenegen.stress_scene(3000, 40, 64, 48, tex_size=32)
with DeviceScene(sc) as ds:
    for _ in range(4):
        st = ds.render(accel=abi.RT_ACCEL_LBVH)
    assert st.graph_replays >= 2
    rng = np.random.default_rng(0)
    o = rng.uniform(-60, 60, (512, 3)).astype(np.float32); d = rng.normal(0, 1, (512, 3)).astype(np.float32)
    sid = rng.integers(1, sc.n_prims + 1, 512).astype(np.int32)
    ds.trace_closest(o, d, 1e-3, accel=abi.RT_ACCEL_LBVH)
    ds.trace_shadow(o, d, sid, 1e-3, None, accel=abi.RT_ACCEL_LBVH8)
with MultiDeviceScene(sc, 1) as ms:
    for _ in range(3):
        ms.render()
    ms.download(); ms.ppm_text()
print("sanitize_case ok")
