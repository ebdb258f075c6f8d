#!/usr/bin/env python
"""LBVH frame (with whatever RT_* knobs the environment sets) == brute-force frame, on two small scenes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from simple_raytracer_b200 import abi, scenegen
from simple_raytracer_b200.renderer import DeviceScene
ok = True
for name, sc in (("coverage", scenegen.coverage_scene(128, 96)), ("stress", scenegen.stress_scene(8000, 100, 320, 240))):
    with DeviceScene(sc) as ds:
        ds.render(accel=abi.RT_ACCEL_BRUTE)
        want = [a.copy() for a in ds.download()]
        for world in (1, 2):
            got = [np.zeros_like(want[0]), np.zeros_like(want[1])]
            for rank in range(world):
                ds.render(accel=abi.RT_ACCEL_LBVH, world_size=world, rank=rank, tile_rows=8)
                ds.download(got[0], got[1])
            same = np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
            ok = ok and same
            print(name, "world", world, "identical" if same else "DIFFERENT")
sys.exit(0 if ok else 1)
