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
print("sanitize_case ok")
