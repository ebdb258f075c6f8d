#!/usr/bin/env python
"""Scratch: where the end-to-end step (upload + render + download) spends its time."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from simple_raytracer_b200 import scenegen
from simple_raytracer_b200.renderer import DeviceScene
import bench

sc = scenegen.stress_scene(1_000_000, 1000, 3840, 2160)
psc = bench.pin_scene(sc)
import torch
rgb = torch.empty((2160, 3840, 3), dtype=torch.uint8).pin_memory().numpy()
flg = torch.empty((2160, 3840), dtype=torch.uint8).pin_memory().numpy()
for it in range(5):
    t0 = time.perf_counter()
    ds = DeviceScene(psc)
    t1 = time.perf_counter()
    st = ds.render(tile_rows=10)
    t2 = time.perf_counter()
    ds.download(rgb, flg)
    t3 = time.perf_counter()
    ds.close()
    t4 = time.perf_counter()
    print(json.dumps(dict(upload_ms=round((t1 - t0) * 1e3, 2), render_call_ms=round((t2 - t1) * 1e3, 2),
                          device_render_ms=round(st.render_ms, 2), download_ms=round((t3 - t2) * 1e3, 2),
                          close_ms=round((t4 - t3) * 1e3, 2), regrows=st.queue_regrows)), flush=True)
