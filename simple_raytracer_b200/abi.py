"""ctypes mirror of include/rt.h and include/rt_host.h.

The product path is the CUDA library `lib/librt_b200.so`; there is no CPU
fallback: `load_rt()` raises when the library has not been built, and every
render call fails when no CUDA device is present.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.path.join(_HERE, "lib")

RT_OK = 0
RT_E_INVALID, RT_E_CUDA, RT_E_NOMEM, RT_E_STATE, RT_E_PARSE, RT_E_IO, RT_E_LIMIT = -1, -2, -3, -4, -5, -6, -7
RT_ACCEL_AUTO, RT_ACCEL_BRUTE, RT_ACCEL_LBVH, RT_ACCEL_LBVH8 = 0, 1, 2, 3

_ERR_NAMES = {
    RT_E_INVALID: "RT_E_INVALID", RT_E_CUDA: "RT_E_CUDA", RT_E_NOMEM: "RT_E_NOMEM",
    RT_E_STATE: "RT_E_STATE", RT_E_PARSE: "RT_E_PARSE", RT_E_IO: "RT_E_IO", RT_E_LIMIT: "RT_E_LIMIT",
}


class RtError(RuntimeError):
    def __init__(self, code: int, where: str, msg: str):
        super().__init__(f"{where}: {_ERR_NAMES.get(code, code)}: {msg}")
        self.code = code
        self.message = msg


class RtMaterial(C.Structure):
    _fields_ = [("od", C.c_float * 3), ("os", C.c_float * 3), ("ka", C.c_float), ("kd", C.c_float),
                ("ks", C.c_float), ("n", C.c_float), ("opacity", C.c_float), ("eta", C.c_float)]


class RtTexture(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgb", C.POINTER(C.c_uint8))]


class RtSphere(C.Structure):
    _fields_ = [("center", C.c_float * 3), ("radius", C.c_float), ("material", C.c_int32),
                ("texture", C.c_int32), ("id", C.c_int32)]


class RtFace(C.Structure):
    _fields_ = [("v", (C.c_float * 3) * 3), ("vn", (C.c_float * 3) * 3), ("vt", (C.c_float * 2) * 3),
                ("smooth", C.c_int32), ("material", C.c_int32), ("texture", C.c_int32), ("id", C.c_int32)]


class RtLight(C.Structure):
    _fields_ = [("xyz", C.c_float * 3), ("w", C.c_float), ("rgb", C.c_float * 3)]


class RtSceneDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32),
                ("eye", C.c_float * 3), ("viewdir", C.c_float * 3), ("updir", C.c_float * 3),
                ("hfov_deg", C.c_float), ("bkg", C.c_float * 3), ("bkg_eta", C.c_float),
                ("n_materials", C.c_int32), ("materials", C.POINTER(RtMaterial)),
                ("n_textures", C.c_int32), ("textures", C.POINTER(RtTexture)),
                ("n_spheres", C.c_int32), ("spheres", C.POINTER(RtSphere)),
                ("n_faces", C.c_int32), ("faces", C.POINTER(RtFace)),
                ("n_lights", C.c_int32), ("lights", C.POINTER(RtLight)),
                ("camera_normalised", C.c_int32)]


class RtRenderParams(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("epsilon", C.c_float), ("accel", C.c_int32),
                ("rank", C.c_int32), ("world_size", C.c_int32), ("tile_rows", C.c_int32),
                ("instrument", C.c_int32), ("profile", C.c_int32), ("subframes", C.c_int32),
                ("work_sharing", C.c_int32)]


class RtStats(C.Structure):
    _fields_ = [("render_ms", C.c_double), ("nodes_total", C.c_uint64), ("nodes_per_level", C.c_uint64 * 8),
                ("rays_primary", C.c_uint64), ("rays_shadow", C.c_uint64), ("rays_transmit", C.c_uint64),
                ("rays_reflect", C.c_uint64), ("bvh_nodes_visited", C.c_uint64), ("prim_tests", C.c_uint64),
                ("texel_fetches", C.c_uint64), ("accel_used", C.c_int32), ("kernel_launches", C.c_int32),
                ("queue_regrows", C.c_int32), ("reserved", C.c_int32),
                ("level_ms", C.c_double * 8), ("combine_ms", C.c_double), ("shadow_rays_walked", C.c_uint64),
                ("trace_ms", C.c_double * 8), ("shadow_ms", C.c_double * 8), ("shade_ms", C.c_double * 8),
                ("shadow_boxes_tested", C.c_uint64), ("shadow_prim_tests", C.c_uint64),
                ("trace_span_ms", C.c_double * 8), ("trace_drain_ms", C.c_double * 8), ("trace_warp_drain_ms", C.c_double * 8),
                ("shadow_span_ms", C.c_double * 8), ("shadow_drain_ms", C.c_double * 8),
                ("shadow_warp_drain_ms", C.c_double * 8),
                ("trace_drain_steps", C.c_double * 8), ("trace_drain_lanes", C.c_double * 8),
                ("trace_drain_steals", C.c_double * 8), ("shadow_drain_steps", C.c_double * 8),
                ("shadow_drain_lanes", C.c_double * 8), ("shadow_chains_redone", C.c_uint64),
                ("bvh_levels4", C.c_int32), ("bvh_levels8", C.c_int32), ("graph_replays", C.c_int32),
                ("shadow_map_hits", C.c_int32), ("spawn_ms", C.c_double * 8)]

    def as_dict(self) -> dict:
        d = {}
        for name, _ in self._fields_:
            v = getattr(self, name)
            d[name] = list(v) if hasattr(v, "__len__") else v
        return d


# The exported symbols of each library — tests check that every one resolves.
RT_SYMBOLS = [
    "rt_upload_scene", "rt_render", "rt_download_framebuffer", "rt_free_scene", "rt_last_error",
    "rt_abi_version", "rt_tile_buffer_bytes", "rt_render_async", "rt_async_frame_status", "rt_release_cached_memory", "rt_format_ppm_device", "rt_deinterleave_tiles", "rt_set_camera",
    "rt_trace_closest", "rt_trace_shadow", "rt_download_primary_ids", "rt_tile_layout", "rt_format_ppm_noalloc",
    "rt_upload_scene_multi", "rt_render_multi", "rt_download_framebuffer_multi", "rt_format_ppm_multi", "rt_num_gpus_multi",
    "rt_free_scene_multi", "rt_download_hit_histograms",
]
RT_HOST_SYMBOLS = [
    "rt_host_parse_scene_file", "rt_host_free_scene_desc", "rt_host_read_texture_p3", "rt_host_free",
    "rt_host_write_ppm", "rt_host_format_ppm", "rt_host_output_path", "rt_host_last_error",
]

_rt = None
_host = None


def _p(t):
    return C.POINTER(t)


def load_rt() -> C.CDLL:
    """Loads the CUDA render core. Fails loudly when it has not been built."""
    global _rt
    if _rt is not None:
        return _rt
    # RT_B200_LIB: another in-tree build of the same library (tools/: compile-time variants for sweeps)
    path = os.environ.get("RT_B200_LIB") or os.path.join(LIB_DIR, "librt_b200.so")
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build it with `make` (nvcc, sm_100a). "
            "simple_raytracer_b200 has no CPU fallback for the render path.")
    lib = C.CDLL(path)
    vp, i32, f32 = C.c_void_p, C.c_int32, C.c_float
    lib.rt_upload_scene.argtypes = [_p(RtSceneDesc), _p(vp)]
    lib.rt_upload_scene.restype = C.c_int
    lib.rt_render.argtypes = [vp, _p(RtRenderParams), _p(RtStats)]
    lib.rt_render.restype = C.c_int
    lib.rt_download_framebuffer.argtypes = [vp, vp, vp]
    lib.rt_download_framebuffer.restype = C.c_int
    lib.rt_free_scene.argtypes = [vp]
    lib.rt_free_scene.restype = None
    lib.rt_last_error.argtypes = []
    lib.rt_last_error.restype = C.c_char_p
    lib.rt_abi_version.argtypes = []
    lib.rt_abi_version.restype = C.c_int
    lib.rt_tile_buffer_bytes.argtypes = [vp, i32, i32]
    lib.rt_tile_buffer_bytes.restype = C.c_size_t
    lib.rt_tile_layout.argtypes = [i32, i32, i32, i32, _p(i32)]
    lib.rt_tile_layout.restype = C.c_int
    lib.rt_render_async.argtypes = [vp, _p(RtRenderParams), vp, vp]
    lib.rt_render_async.restype = C.c_int
    lib.rt_format_ppm_device.argtypes = [vp, vp, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
    lib.rt_format_ppm_device.restype = C.c_int
    lib.rt_format_ppm_noalloc.argtypes = [vp, i32, i32, _p(f32), vp, C.c_size_t, vp, vp]
    lib.rt_format_ppm_noalloc.restype = C.c_int
    lib.rt_upload_scene_multi.argtypes = [_p(RtSceneDesc), i32, _p(i32), _p(vp)]
    lib.rt_upload_scene_multi.restype = C.c_int
    lib.rt_render_multi.argtypes = [vp, _p(RtRenderParams), _p(RtStats)]
    lib.rt_render_multi.restype = C.c_int
    lib.rt_download_framebuffer_multi.argtypes = [vp, vp, vp]
    lib.rt_download_framebuffer_multi.restype = C.c_int
    lib.rt_format_ppm_multi.argtypes = [vp, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
    lib.rt_format_ppm_multi.restype = C.c_int
    lib.rt_num_gpus_multi.argtypes = [vp]
    lib.rt_num_gpus_multi.restype = C.c_int
    lib.rt_free_scene_multi.argtypes = [vp]
    lib.rt_free_scene_multi.restype = None
    lib.rt_download_hit_histograms.argtypes = [vp, vp, vp]
    lib.rt_download_hit_histograms.restype = C.c_int
    lib.rt_release_cached_memory.argtypes = []
    lib.rt_release_cached_memory.restype = C.c_size_t
    lib.rt_async_frame_status.argtypes = [vp]
    lib.rt_async_frame_status.restype = C.c_int
    lib.rt_deinterleave_tiles.argtypes = [vp, vp, vp, i32, i32, vp]
    lib.rt_deinterleave_tiles.restype = C.c_int
    lib.rt_set_camera.argtypes = [vp, _p(f32), _p(f32), _p(f32), f32]
    lib.rt_set_camera.restype = C.c_int
    lib.rt_trace_closest.argtypes = [vp, i32, i32, vp, vp, f32, vp, vp, vp, vp]
    lib.rt_trace_closest.restype = C.c_int
    lib.rt_trace_shadow.argtypes = [vp, i32, i32, vp, vp, vp, f32, vp, vp, vp]
    lib.rt_trace_shadow.restype = C.c_int
    lib.rt_download_primary_ids.argtypes = [vp, vp]
    lib.rt_download_primary_ids.restype = C.c_int
    _rt = lib
    return lib


def load_host() -> C.CDLL:
    """Loads the CPU front-end (scene parser, P3 reader / writer)."""
    global _host
    if _host is not None:
        return _host
    path = os.path.join(LIB_DIR, "librt_host.so")
    if not os.path.exists(path):
        raise ImportError(f"{path} is missing: build it with `make`.")
    lib = C.CDLL(path)
    vp, i32 = C.c_void_p, C.c_int32
    lib.rt_host_parse_scene_file.argtypes = [C.c_char_p, C.c_char_p, _p(_p(RtSceneDesc))]
    lib.rt_host_parse_scene_file.restype = C.c_int
    lib.rt_host_free_scene_desc.argtypes = [_p(RtSceneDesc)]
    lib.rt_host_free_scene_desc.restype = None
    lib.rt_host_read_texture_p3.argtypes = [C.c_char_p, _p(i32), _p(i32), _p(_p(C.c_uint8))]
    lib.rt_host_read_texture_p3.restype = C.c_int
    lib.rt_host_free.argtypes = [vp]
    lib.rt_host_free.restype = None
    lib.rt_host_write_ppm.argtypes = [C.c_char_p, i32, i32, vp, vp, _p(C.c_float)]
    lib.rt_host_write_ppm.restype = C.c_int
    lib.rt_host_format_ppm.argtypes = [i32, i32, vp, vp, _p(C.c_float), _p(vp), _p(C.c_size_t)]
    lib.rt_host_format_ppm.restype = C.c_int
    lib.rt_host_output_path.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t]
    lib.rt_host_output_path.restype = C.c_int
    lib.rt_host_last_error.argtypes = []
    lib.rt_host_last_error.restype = C.c_char_p
    _host = lib
    return lib


def check(rc: int, where: str, host: bool = False) -> None:
    if rc == RT_OK:
        return
    lib = load_host() if host else load_rt()
    msg = (lib.rt_host_last_error() if host else lib.rt_last_error()) or b""
    raise RtError(rc, where, msg.decode("utf-8", "replace"))
