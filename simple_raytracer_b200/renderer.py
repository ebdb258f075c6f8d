"""Host-side mirror of the render seam: `DeviceScene` wraps the three C-ABI calls
that replace create_view_window_and_ray_trace (main.cpp:607):

    rt_upload_scene -> rt_render -> rt_download_framebuffer

Everything here goes through lib/librt_b200.so; there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np

from . import abi
from .scene import Scene


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_params(max_depth: int = 4, epsilon: float = 1.0e-3, accel: int = abi.RT_ACCEL_AUTO, rank: int = 0,
                world_size: int = 1, tile_rows: int = 8, instrument: bool = False, profile: bool = False,
                subframes: int = 0, work_sharing: int = 0) -> abi.RtRenderParams:
    return abi.RtRenderParams(max_depth, epsilon, accel, rank, world_size, tile_rows, 1 if instrument else 0,
                              1 if profile else 0, subframes, work_sharing)


class DeviceScene:
    """A scene resident on the current CUDA device."""

    def __init__(self, scene: Scene):
        self._lib = abi.load_rt()
        self.scene = scene
        self._h = C.c_void_p()
        rc = self._lib.rt_upload_scene(C.byref(scene.desc()), C.byref(self._h))
        abi.check(rc, "rt_upload_scene")

    # -- lifetime ---------------------------------------------------------------
    def close(self) -> None:
        if self._h:
            self._lib.rt_free_scene(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self) -> C.c_void_p:
        return self._h

    # -- the hot path ---------------------------------------------------------------
    def render(self, params: Optional[abi.RtRenderParams] = None, **kw) -> abi.RtStats:
        """Synchronous render of this rank's row tiles (rt_render)."""
        p = params if params is not None else make_params(**kw)
        st = abi.RtStats()
        abi.check(self._lib.rt_render(self._h, C.byref(p), C.byref(st)), "rt_render")
        return st

    def render_async(self, params: abi.RtRenderParams, dev_tiles_ptr: int = 0, stream_ptr: int = 0) -> None:
        """Enqueue-only render into a caller-provided device tile buffer (rt_render_async)."""
        abi.check(self._lib.rt_render_async(self._h, C.byref(params), C.c_void_p(dev_tiles_ptr or None),
                                            C.c_void_p(stream_ptr or None)), "rt_render_async")

    def async_frame_ok(self) -> bool:
        """After the async frame's stream was synchronised: False = a queue overflowed (queues
        have been grown, render the frame again) (rt_async_frame_status)."""
        rc = self._lib.rt_async_frame_status(self._h)
        if rc == abi.RT_E_NOMEM:
            return False
        abi.check(rc, "rt_async_frame_status")
        return True

    def download(self, rgb: Optional[np.ndarray] = None, flags: Optional[np.ndarray] = None
                 ) -> Tuple[np.ndarray, np.ndarray]:
        """-> (rgb uint8 [H,W,3], flags uint8 [H,W]); see rt_download_framebuffer."""
        H, W = self.scene.height, self.scene.width
        if rgb is None:
            rgb = np.zeros((H, W, 3), dtype=np.uint8)
        if flags is None:
            flags = np.zeros((H, W), dtype=np.uint8)
        abi.check(self._lib.rt_download_framebuffer(self._h, _ptr(rgb), _ptr(flags)), "rt_download_framebuffer")
        return rgb, flags

    def ppm_text(self, dev_rgba_ptr: int = 0) -> bytes:
        """The P3 file of the frame, formatted on the device (rt_format_ppm_device)."""
        text = C.c_void_p(); n = C.c_size_t()
        abi.check(self._lib.rt_format_ppm_device(self._h, C.c_void_p(dev_rgba_ptr or None), C.byref(text), C.byref(n)),
                  "rt_format_ppm_device")
        return C.string_at(text, n.value)

    def download_primary_ids(self) -> np.ndarray:
        ids = np.zeros((self.scene.height, self.scene.width), dtype=np.int32)
        abi.check(self._lib.rt_download_primary_ids(self._h, _ptr(ids)), "rt_download_primary_ids")
        return ids

    def hit_histograms(self) -> Tuple[np.ndarray, np.ndarray]:
        """-> (hits uint64 [8][n_objects], shadow uint64 [8][n_lights][3]) of the last LBVH frame (rt_download_hit_histograms)."""
        hits = np.zeros((8, self.scene.n_prims), dtype=np.uint64)
        shadow = np.zeros((8, max(len(self.scene.lights), 1), 3), dtype=np.uint64)
        abi.check(self._lib.rt_download_hit_histograms(self._h, _ptr(hits), _ptr(shadow)), "rt_download_hit_histograms")
        return hits, shadow[:, :len(self.scene.lights)]

    def tile_buffer_bytes(self, world_size: int, tile_rows: int) -> int:
        return int(self._lib.rt_tile_buffer_bytes(self._h, world_size, tile_rows))

    def deinterleave(self, gathered_ptr: int, out_ptr: int, world_size: int, tile_rows: int, stream_ptr: int = 0) -> None:
        abi.check(self._lib.rt_deinterleave_tiles(self._h, C.c_void_p(gathered_ptr), C.c_void_p(out_ptr), world_size,
                                                  tile_rows, C.c_void_p(stream_ptr or None)), "rt_deinterleave_tiles")

    def set_camera(self, eye, viewdir, updir, hfov_deg: float) -> None:
        f3 = C.c_float * 3
        abi.check(self._lib.rt_set_camera(self._h, f3(*map(float, eye)), f3(*map(float, viewdir)),
                                          f3(*map(float, updir)), float(hfov_deg)), "rt_set_camera")

    # -- parity probes ----------------------------------------------------------------
    def trace_closest(self, origins, dirs, tmin: float, accel: int = abi.RT_ACCEL_AUTO):
        o = np.ascontiguousarray(origins, dtype=np.float32)
        d = np.ascontiguousarray(dirs, dtype=np.float32)
        n = len(o)
        ids = np.zeros(n, np.int32); t = np.zeros(n, np.float32)
        nrm = np.zeros((n, 3), np.float32); bary = np.zeros((n, 3), np.float32)
        abi.check(self._lib.rt_trace_closest(self._h, accel, n, _ptr(o), _ptr(d), tmin, _ptr(ids), _ptr(t), _ptr(nrm),
                                             _ptr(bary)), "rt_trace_closest")
        return ids, t, nrm, bary

    def trace_shadow(self, origins, dirs, self_id, tmin: float, tmax=None, accel: int = abi.RT_ACCEL_AUTO):
        o = np.ascontiguousarray(origins, dtype=np.float32)
        d = np.ascontiguousarray(dirs, dtype=np.float32)
        sid = np.ascontiguousarray(self_id, dtype=np.int32)
        tm = None if tmax is None else np.ascontiguousarray(tmax, dtype=np.float32)
        n = len(o)
        mask = np.zeros(n, np.float32); hits = np.zeros(n, np.int32)
        abi.check(self._lib.rt_trace_shadow(self._h, accel, n, _ptr(o), _ptr(d), _ptr(sid), tmin, _ptr(tm), _ptr(mask),
                                            _ptr(hits)), "rt_trace_shadow")
        return mask, hits


def render_scene(scene: Scene, **kw) -> Tuple[np.ndarray, np.ndarray, abi.RtStats]:
    """One-shot: upload, render, download. -> (rgb [H,W,3], flags [H,W], stats)"""
    with DeviceScene(scene) as ds:
        st = ds.render(**kw)
        rgb, flags = ds.download()
    return rgb, flags, st


class MultiDeviceScene:
    """A scene resident on N GPUs of this node, driven by ONE process through rt_*_multi
    (NCCL communicator, broadcast of the built scene, allgather of the tiles inside the library)."""

    def __init__(self, scene: Scene, num_gpus: int, device_ids=None):
        self._lib = abi.load_rt()
        self.scene = scene
        self._h = C.c_void_p()
        ids = None if device_ids is None else (C.c_int32 * num_gpus)(*device_ids)
        abi.check(self._lib.rt_upload_scene_multi(C.byref(scene.desc()), num_gpus, ids, C.byref(self._h)),
                  "rt_upload_scene_multi")

    def close(self) -> None:
        if self._h:
            self._lib.rt_free_scene_multi(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def num_gpus(self) -> int:
        return int(self._lib.rt_num_gpus_multi(self._h))

    def render(self, params: Optional[abi.RtRenderParams] = None, **kw) -> abi.RtStats:
        kw.setdefault("tile_rows", 0)
        p = params if params is not None else make_params(**kw)
        st = abi.RtStats()
        abi.check(self._lib.rt_render_multi(self._h, C.byref(p), C.byref(st)), "rt_render_multi")
        return st

    def download(self) -> Tuple[np.ndarray, np.ndarray]:
        H, W = self.scene.height, self.scene.width
        rgb = np.zeros((H, W, 3), dtype=np.uint8)
        flags = np.zeros((H, W), dtype=np.uint8)
        abi.check(self._lib.rt_download_framebuffer_multi(self._h, _ptr(rgb), _ptr(flags)), "rt_download_framebuffer_multi")
        return rgb, flags

    def ppm_text(self) -> bytes:
        text = C.c_void_p(); n = C.c_size_t()
        abi.check(self._lib.rt_format_ppm_multi(self._h, C.byref(text), C.byref(n)), "rt_format_ppm_multi")
        return C.string_at(text, n.value)
