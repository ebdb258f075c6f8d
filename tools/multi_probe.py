#!/usr/bin/env python
"""One process, N GPUs through rt_*_multi: upload (H2D once + pack + LBVH build + NCCL broadcast), frames, download.
usage: multi_probe.py N [width height]"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from simple_raytracer_b200 import scenegen
from simple_raytracer_b200.renderer import DeviceScene, MultiDeviceScene

n = int(sys.argv[1]); w = int(sys.argv[2]) if len(sys.argv) > 2 else 3840; h = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
sc = scenegen.stress_scene(1_000_000, 1000, w, h)
out = {"gpus": n, "width": w, "height": h}
for rep in range(2):                      # second round: warm allocator cache, NCCL communicators exist
    t0 = time.perf_counter()
    ms = MultiDeviceScene(sc, n)
    t1 = time.perf_counter()
    frames = []
    for k in range(8):
        st = ms.render()
        frames.append(round(st.render_ms, 3))
    t2 = time.perf_counter()
    rgb, fl = ms.download()
    t3 = time.perf_counter()
    ms.close()
    out[f"round{rep}"] = {"upload_ms": round((t1 - t0) * 1e3, 2), "first_frame_ms": frames[0], "frames_ms": frames,
                          "download_ms": round((t3 - t2) * 1e3, 2), "graph_replays": int(st.graph_replays)}
import hashlib
out["sha256_rgb"] = hashlib.sha256(rgb.tobytes()).hexdigest()
with DeviceScene(sc) as ds:
    ds.render(tile_rows=10)
    r1, f1 = ds.download()
out["equals_single_gpu"] = bool(np.array_equal(r1, rgb) and np.array_equal(f1, fl))
print(json.dumps(out), flush=True)
