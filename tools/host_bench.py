#!/usr/bin/env python
"""CPU timing of the host front-end rows next to the render core:
P3 frame writer, P3 texture loader, scene parser.  No GPU needed."""
import ctypes as C, json, os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from simple_raytracer_b200 import abi, scenegen
from simple_raytracer_b200.scene import Scene

host = abi.load_host()
out = {}
rng = np.random.default_rng(1)
# f-1: 4K frame -> P3 text
W, H = 3840, 2160
rgb = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
flags = np.zeros((H, W), dtype=np.uint8)
bkg = (C.c_float * 3)(0.2, 0.3, 0.5)
txt = C.c_char_p(); n = C.c_size_t()
host.rt_host_format_ppm.argtypes = [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t)]
best = 1e9
for _ in range(3):
    t0 = time.perf_counter()
    rc = host.rt_host_format_ppm(W, H, rgb.ctypes.data, flags.ctypes.data, bkg, C.byref(txt), C.byref(n))
    best = min(best, time.perf_counter() - t0)
    assert rc == 0
    host.rt_host_free(txt)
out["format_ppm_4k_ms"] = round(best * 1e3, 1)
out["format_ppm_bytes"] = n.value
out["format_ppm_GBps"] = round(n.value / best / 1e9, 2)
with tempfile.TemporaryDirectory() as td:
    # f-2: 2048x1024 P3 texture
    tex = scenegen.make_texture("earth", 2048, 1024, 7)
    p = os.path.join(td, "t.ppm")
    open(p, "w").write(scenegen.p3_text(tex))
    sz = os.path.getsize(p)
    w_ = C.c_int32(); h_ = C.c_int32(); buf = C.POINTER(C.c_uint8)()
    for thr in ("1", ""):
        if thr:
            os.environ["RT_HOST_THREADS"] = thr
        else:
            os.environ.pop("RT_HOST_THREADS", None)
        best = 1e9
        for _ in range(4):
            t0 = time.perf_counter()
            rc = host.rt_host_read_texture_p3(os.fsencode(p), C.byref(w_), C.byref(h_), C.byref(buf))
            best = min(best, time.perf_counter() - t0)
            assert rc == 0
            host.rt_host_free(buf)
        tag = "1thread" if thr else "allthreads"
        out[f"read_p3_2048x1024_ms_{tag}"] = round(best * 1e3, 1)
        out[f"read_p3_MBps_{tag}"] = round(sz / best / 1e6, 1)
    # f-3: scene text of a triangle soup
    ntri = int(os.environ.get("HOST_BENCH_TRIS", "200000"))
    sc = scenegen.stress_scene(ntri, 1000, 64, 64, tex_size=16)
    paths = []
    for k, t in enumerate(sc.textures):
        tp = os.path.join(td, f"tex{k}.ppm"); open(tp, "w").write(scenegen.p3_text(t)); paths.append(tp)
    sp = os.path.join(td, "scene.txt")
    t0 = time.perf_counter(); open(sp, "w").write(sc.to_text(paths)); gen = time.perf_counter() - t0
    sz = os.path.getsize(sp)
    out["parse_scene_tris"] = ntri
    desc = C.POINTER(abi.RtSceneDesc)()
    for thr in ("1", ""):
        if thr:
            os.environ["RT_HOST_THREADS"] = thr
        else:
            os.environ.pop("RT_HOST_THREADS", None)
        best = 1e9
        for _ in range(3):
            t0 = time.perf_counter()
            rc = host.rt_host_parse_scene_file(os.fsencode(sp), None, C.byref(desc))
            best = min(best, time.perf_counter() - t0)
            assert rc == 0 and desc.contents.n_faces == ntri
            host.rt_host_free_scene_desc(desc)
        tag = "1thread" if thr else "allthreads"
        out[f"parse_scene_ms_{tag}"] = round(best * 1e3, 1)
        out[f"parse_scene_MBps_{tag}"] = round(sz / best / 1e6, 1)
    out["host_cpus"] = os.cpu_count()
    out["scene_file_MB"] = round(sz / 1e6, 1)
print(json.dumps(out))
