"""P3 texture reader and P3 frame writer of the host front-end (include/rt_host.h):
utility.h:59-139 and main.cpp:613-650 of the reference."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

from . import abi


def read_texture_p3(path: str) -> np.ndarray:
    """-> uint8 [height][width][3]"""
    host = abi.load_host()
    w, h = C.c_int32(), C.c_int32()
    px = C.POINTER(C.c_uint8)()
    abi.check(host.rt_host_read_texture_p3(os.fsencode(path), C.byref(w), C.byref(h), C.byref(px)),
              f"rt_host_read_texture_p3({path})", host=True)
    try:
        return np.frombuffer(C.string_at(px, w.value * h.value * 3), dtype=np.uint8).reshape(h.value, w.value, 3).copy()
    finally:
        host.rt_host_free(px)


def _args(rgb: np.ndarray, flags: Optional[np.ndarray], bkg: Optional[Sequence[float]]):
    rgb = np.ascontiguousarray(rgb, dtype=np.uint8)
    h, w, _ = rgb.shape
    fl = None if flags is None else np.ascontiguousarray(flags, dtype=np.uint8)
    b = None if bkg is None else (C.c_float * 3)(*[float(x) for x in bkg])
    return rgb, fl, b, w, h


def format_ppm(rgb: np.ndarray, flags: Optional[np.ndarray] = None, bkg: Optional[Sequence[float]] = None) -> str:
    host = abi.load_host()
    rgb, fl, b, w, h = _args(rgb, flags, bkg)
    text, n = C.c_void_p(), C.c_size_t()
    abi.check(host.rt_host_format_ppm(w, h, rgb.ctypes.data_as(C.c_void_p), None if fl is None else fl.ctypes.data_as(C.c_void_p),
                                      b, C.byref(text), C.byref(n)), "rt_host_format_ppm", host=True)
    try:
        return C.string_at(text, n.value).decode("ascii")
    finally:
        host.rt_host_free(text)


def write_ppm(path: str, rgb: np.ndarray, flags: Optional[np.ndarray] = None, bkg: Optional[Sequence[float]] = None) -> None:
    host = abi.load_host()
    rgb, fl, b, w, h = _args(rgb, flags, bkg)
    abi.check(host.rt_host_write_ppm(os.fsencode(path), w, h, rgb.ctypes.data_as(C.c_void_p),
                                     None if fl is None else fl.ctypes.data_as(C.c_void_p), b), "rt_host_write_ppm", host=True)


def output_path(scene_path: str) -> str:
    host = abi.load_host()
    buf = C.create_string_buffer(4096)
    abi.check(host.rt_host_output_path(os.fsencode(scene_path), buf, 4096), "rt_host_output_path", host=True)
    return buf.value.decode()
