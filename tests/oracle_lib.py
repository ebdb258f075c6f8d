"""ctypes bindings of the TEST ORACLE (oracle/liboracle.so) and of the reference
itself compiled from /root/reference (oracle/_ref/libref.so).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module; the product never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from simple_raytracer_b200 import abi
from simple_raytracer_b200.scene import Scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
RTO_NAN = -2147483648


class RtoParams(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("epsilon", C.c_float)]


class RtoStats(C.Structure):
    _fields_ = [("nodes_total", C.c_uint64), ("nodes_per_level", C.c_uint64 * 8),
                ("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_transmit", C.c_uint64),
                ("rays_reflect", C.c_uint64), ("prim_tests", C.c_uint64), ("texel_fetches", C.c_uint64)]

    @property
    def rays_total(self) -> int:
        return self.rays_primary + self.rays_shadow + self.rays_transmit + self.rays_reflect


_oracle = None
_ref = None


def build_oracle() -> None:
    subprocess.run(["make", "-C", ORACLE_DIR, "all"], check=True, capture_output=True)


def load_oracle() -> C.CDLL:
    global _oracle
    if _oracle is None:
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            build_oracle()
        lib = C.CDLL(path)
        vp = C.c_void_p
        lib.rto_render.argtypes = [C.POINTER(abi.RtSceneDesc), C.POINTER(RtoParams), C.c_int64, vp, vp, vp,
                                   C.POINTER(RtoStats)]
        lib.rto_render.restype = C.c_int
        lib.rto_primary_rays.argtypes = [C.POINTER(abi.RtSceneDesc), C.c_int64, vp, vp, vp]
        lib.rto_primary_rays.restype = C.c_int
        lib.rto_trace_closest.argtypes = [C.POINTER(abi.RtSceneDesc), C.c_int32, vp, vp, C.c_float, vp, vp, vp, vp]
        lib.rto_trace_closest.restype = C.c_int
        lib.rto_trace_shadow.argtypes = [C.POINTER(abi.RtSceneDesc), C.c_int32, vp, vp, vp, C.c_float, vp, vp, vp]
        lib.rto_trace_shadow.restype = C.c_int
        _oracle = lib
    return _oracle


def have_ref() -> bool:
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libref.so"))


def load_ref() -> C.CDLL:
    global _ref
    if _ref is None:
        lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", "libref.so"))
        vp = C.c_void_p
        lib.ref_render.argtypes = [C.POINTER(abi.RtSceneDesc), C.c_int, C.c_float, vp, C.POINTER(C.c_double)]
        lib.ref_render.restype = C.c_int
        lib.ref_render_pixels.argtypes = [C.POINTER(abi.RtSceneDesc), C.c_int, C.c_float, C.c_int64, vp, vp, vp,
                                          C.POINTER(C.c_double)]
        lib.ref_render_pixels.restype = C.c_int
        lib.ref_trace_closest.argtypes = [C.POINTER(abi.RtSceneDesc), C.c_int32, vp, vp, C.c_float, vp, vp, vp, vp]
        lib.ref_trace_closest.restype = C.c_int
        lib.ref_trace_shadow.argtypes = [C.POINTER(abi.RtSceneDesc), C.c_int32, vp, vp, vp, C.c_float, vp, vp, vp]
        lib.ref_trace_shadow.restype = C.c_int
        _ref = lib
    return _ref


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _pixels(scene: Scene, pixels):
    if pixels is None:
        return scene.n_pixels, None
    px = np.ascontiguousarray(np.asarray(pixels, dtype=np.int64))
    return len(px), px


# ---- the plain-C restatement ------------------------------------------------
def oracle_render(scene: Scene, max_depth: int = 4, epsilon: float = 1.0e-3, pixels=None):
    """-> (raw int32 [n,3] exactly as the reference's matt(i,j,c), primary ids [n], RtoStats)"""
    lib = load_oracle()
    n, px = _pixels(scene, pixels)
    raw = np.empty((n, 3), dtype=np.int32)
    ids = np.empty(n, dtype=np.int32)
    st = RtoStats()
    prm = RtoParams(max_depth, epsilon)
    rc = lib.rto_render(C.byref(scene.desc()), C.byref(prm), n, _ptr(px), _ptr(raw), _ptr(ids), C.byref(st))
    if rc != 0:
        raise RuntimeError(f"rto_render failed: {rc}")
    return raw, ids, st


def oracle_hit_histograms(scene: Scene, max_depth: int = 4, epsilon: float = 1.0e-3):
    """-> (hits uint64 [8][n_objects], shadow uint64 [8][n_lights][3]) of a whole-frame oracle render:
    which objects ShadeRay was invoked on per generation, and the class of the carried shadow_mask after
    each light (== 0 / in between / == 1); see oracle/rt_oracle.c: rto_set_histograms."""
    lib = load_oracle()
    lib.rto_set_histograms.argtypes = [C.c_void_p, C.c_void_p]
    lib.rto_set_histograms.restype = None
    hits = np.zeros((8, scene.n_prims), dtype=np.uint64)
    shadow = np.zeros((8, max(len(scene.lights), 1), 3), dtype=np.uint64)
    lib.rto_set_histograms(_ptr(hits), _ptr(shadow))
    try:
        oracle_render(scene, max_depth, epsilon)
    finally:
        lib.rto_set_histograms(None, None)
    return hits, shadow[:, :len(scene.lights)]


def primary_rays(scene: Scene, rows):
    """Primary rays of whole image rows -> (origins [n,3], dirs [n,3]) float32."""
    lib = load_oracle()
    px = np.concatenate([np.arange(scene.width, dtype=np.int64) + int(r) * scene.width for r in rows])
    o = np.zeros((len(px), 3), np.float32); d = np.zeros((len(px), 3), np.float32)
    rc = lib.rto_primary_rays(C.byref(scene.desc()), len(px), _ptr(px), _ptr(o), _ptr(d))
    assert rc == 0
    return o, d


def oracle_trace_closest(scene: Scene, origins, dirs, tmin: float):
    lib = load_oracle()
    o = np.ascontiguousarray(origins, dtype=np.float32)
    d = np.ascontiguousarray(dirs, dtype=np.float32)
    n = len(o)
    ids = np.zeros(n, np.int32); t = np.zeros(n, np.float32)
    nrm = np.zeros((n, 3), np.float32); bary = np.zeros((n, 3), np.float32)
    rc = lib.rto_trace_closest(C.byref(scene.desc()), n, _ptr(o), _ptr(d), tmin, _ptr(ids), _ptr(t), _ptr(nrm), _ptr(bary))
    assert rc == 0
    return ids, t, nrm, bary


def oracle_trace_shadow(scene: Scene, origins, dirs, self_id, tmin: float, tmax=None):
    lib = load_oracle()
    o = np.ascontiguousarray(origins, dtype=np.float32)
    d = np.ascontiguousarray(dirs, dtype=np.float32)
    sid = np.ascontiguousarray(self_id, dtype=np.int32)
    tm = None if tmax is None else np.ascontiguousarray(tmax, dtype=np.float32)
    n = len(o)
    mask = np.zeros(n, np.float32); hits = np.zeros(n, np.int32)
    rc = lib.rto_trace_shadow(C.byref(scene.desc()), n, _ptr(o), _ptr(d), _ptr(sid), tmin, _ptr(tm), _ptr(mask), _ptr(hits))
    assert rc == 0
    return mask, hits


# ---- the reference itself -----------------------------------------------------
def ref_render(scene: Scene, max_depth: int = 4, epsilon: float = 1.0e-3):
    """Whole frame through the reference's create_view_window_and_ray_trace.
    -> (raw int32 [H*W,3], seconds spent inside that call)"""
    lib = load_ref()
    raw = np.empty((scene.n_pixels, 3), dtype=np.int32)
    secs = C.c_double(0.0)
    rc = lib.ref_render(C.byref(scene.desc()), max_depth, epsilon, _ptr(raw), C.byref(secs))
    assert rc == 0
    return raw, secs.value


def ref_render_pixels(scene: Scene, pixels, max_depth: int = 4, epsilon: float = 1.0e-3):
    lib = load_ref()
    n, px = _pixels(scene, pixels)
    raw = np.empty((n, 3), dtype=np.int32)
    ids = np.empty(n, dtype=np.int32)
    secs = C.c_double(0.0)
    rc = lib.ref_render_pixels(C.byref(scene.desc()), max_depth, epsilon, n, _ptr(px), _ptr(raw), _ptr(ids), C.byref(secs))
    assert rc == 0
    return raw, ids, secs.value


def ref_trace_closest(scene: Scene, origins, dirs, tmin: float):
    lib = load_ref()
    o = np.ascontiguousarray(origins, dtype=np.float32)
    d = np.ascontiguousarray(dirs, dtype=np.float32)
    n = len(o)
    ids = np.zeros(n, np.int32); t = np.zeros(n, np.float32)
    nrm = np.zeros((n, 3), np.float32); bary = np.zeros((n, 3), np.float32)
    rc = lib.ref_trace_closest(C.byref(scene.desc()), n, _ptr(o), _ptr(d), tmin, _ptr(ids), _ptr(t), _ptr(nrm), _ptr(bary))
    assert rc == 0
    return ids, t, nrm, bary


def ref_trace_shadow(scene: Scene, origins, dirs, self_id, tmin: float, tmax=None):
    lib = load_ref()
    o = np.ascontiguousarray(origins, dtype=np.float32)
    d = np.ascontiguousarray(dirs, dtype=np.float32)
    sid = np.ascontiguousarray(self_id, dtype=np.int32)
    tm = None if tmax is None else np.ascontiguousarray(tmax, dtype=np.float32)
    n = len(o)
    mask = np.zeros(n, np.float32); hits = np.zeros(n, np.int32)
    rc = lib.ref_trace_shadow(C.byref(scene.desc()), n, _ptr(o), _ptr(d), _ptr(sid), tmin, _ptr(tm), _ptr(mask), _ptr(hits))
    assert rc == 0
    return mask, hits


# ---- helpers -------------------------------------------------------------------
def raw_to_rgb8(raw: np.ndarray, primary_ids=None):
    """Splits raw ints into (rgb uint8 [n,3], flags uint8 [n]) = rt_download_framebuffer's layout."""
    nan = raw == RTO_NAN
    rgb = np.clip(np.where(nan, 0, raw), 0, 255).astype(np.uint8)
    flags = (nan[:, 0] * 1 + nan[:, 1] * 2 + nan[:, 2] * 4).astype(np.uint8)
    if primary_ids is not None:
        flags = flags | ((np.asarray(primary_ids) == 0) * 8).astype(np.uint8)
    return rgb, flags
