#!/usr/bin/env python
"""Device P3 formatter vs the host one on a 4K frame (timing + identity)."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simple_raytracer_b200 import scenegen, ppm
from simple_raytracer_b200.renderer import DeviceScene
sc = scenegen.stress_scene(100000, 100, 3840, 2160)
with DeviceScene(sc) as ds:
    ds.render()
    rgb, flags = ds.download()
    t = []
    for _ in range(4):
        t0 = time.perf_counter(); txt = ds.ppm_text(); t.append(time.perf_counter() - t0)
    import ctypes as C
    from simple_raytracer_b200 import abi
    raw = []
    for _ in range(4):
        text = C.c_void_p(); n = C.c_size_t()
        t0 = time.perf_counter()
        rc = ds._lib.rt_format_ppm_device(ds._h, None, C.byref(text), C.byref(n))
        raw.append(time.perf_counter() - t0)
        assert rc == 0
    t0 = time.perf_counter(); want = ppm.format_ppm(rgb, flags, bkg=sc.bkg); host = time.perf_counter() - t0
    want = want if isinstance(want, bytes) else want.encode()
    print(json.dumps(dict(c_call_ms=[round(x * 1e3, 2) for x in raw], device_ms=[round(x * 1e3, 2) for x in t], host_ms=round(host * 1e3, 1), bytes=len(txt), same=txt == want)))
