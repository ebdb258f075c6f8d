"""rt_format_ppm_noalloc (include/rt.h) against rt_format_ppm_device.

The entry point has no counterpart in the reference: it was added
for callers that cannot call cudaMalloc.  It formats a frame that is already on the
device into a caller-provided device buffer, using a static __device__ scratch array instead of
workspace allocations.  The bar is the one the other P3 tests use: the text is byte-identical.

Collected last on purpose.  Passes alone in a fresh process
and after the rest of the suite; the two calls in a row go from a larger frame to a smaller one
and back, so a call that relied on what the previous one left in the shared scratch would show.
"""
import ctypes as C

import numpy as np
import pytest

from simple_raytracer_b200 import abi, scenegen
from simple_raytracer_b200.renderer import DeviceScene, make_params
from tests import common

pytestmark = pytest.mark.gpu


def _both_texts(scene, torch):
    """-> (text of rt_format_ppm_device, text of rt_format_ppm_noalloc, length it reported)"""
    lib = abi.load_rt()
    n = scene.width * scene.height
    with DeviceScene(scene) as ds:
        # the frame in caller-owned device memory: world_size 1 -> the tile buffer IS scanline order
        frame = torch.zeros(ds.tile_buffer_bytes(1, 8), dtype=torch.uint8, device="cuda")
        ds.render(make_params())                       # sizes the queues
        ds.render_async(make_params(), frame.data_ptr(), 0)
        torch.cuda.synchronize()
        assert ds.async_frame_ok()
        want = ds.ppm_text(frame.data_ptr())
    cap = 41 + 64 * n
    text = torch.full((cap,), 0x23, dtype=torch.uint8, device="cuda")      # '#': never part of a P3 file
    length = torch.zeros(1, dtype=torch.int64, device="cuda")
    bkg = (C.c_float * 3)(*[float(x) for x in scene.bkg])
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.rt_format_ppm_noalloc(C.c_void_p(frame.data_ptr()), scene.width, scene.height, bkg,
                                   C.c_void_p(text.data_ptr()), cap, C.c_void_p(length.data_ptr()), C.c_void_p(stream))
    abi.check(rc, "rt_format_ppm_noalloc")
    torch.cuda.synchronize()
    got_len = int(length.item())
    host = text.cpu().numpy()
    assert got_len <= cap
    assert (host[got_len:] == 0x23).all()              # nothing written past the reported length
    return want, host[:got_len].tobytes(), got_len


def test_no_alloc_formatter_equals_the_allocating_one_twice_in_a_row():
    torch = pytest.importorskip("torch")
    test1, _ = common.golden_frame("reflection_transparency_tests/Test1")          # NaN pixels
    cases = [scenegen.coverage_scene(131, 77).copy(bkg=(2.0, -0.5, 0.25)),         # out-of-range bkg, ragged last block
             test1,
             scenegen.coverage_scene(16, 2),                                        # fewer pixels than one block
             scenegen.coverage_scene(131, 77).copy(bkg=(2.0, -0.5, 0.25))]          # ... and the first one again
    texts = []
    for sc in cases:
        want, got, got_len = _both_texts(sc, torch)
        assert got_len == len(want), (sc.width, sc.height, got_len, len(want))
        assert got == want, (sc.width, sc.height)
        texts.append(got)
    assert texts[0] == texts[3]
    assert b"18446744071562067968" in texts[1]         # the NaN token really was exercised


def test_no_alloc_formatter_reports_a_short_buffer_and_refuses_oversize_frames():
    torch = pytest.importorskip("torch")
    lib = abi.load_rt()
    w, h = 64, 33
    frame = torch.randint(0, 256, (w * h * 4,), dtype=torch.uint8, device="cuda")
    frame.view(-1, 4)[:, 3] = 0                         # no flags: plain 0..255 values
    bkg = (C.c_float * 3)(0.1, 0.2, 0.3)
    length = torch.zeros(1, dtype=torch.int64, device="cuda")
    cap = 4000                                          # far too small for 2112 lines
    text = torch.full((cap + 64,), 0x23, dtype=torch.uint8, device="cuda")
    rc = lib.rt_format_ppm_noalloc(C.c_void_p(frame.data_ptr()), w, h, bkg, C.c_void_p(text.data_ptr()), cap,
                                   C.c_void_p(length.data_ptr()), None)
    abi.check(rc, "rt_format_ppm_noalloc")
    torch.cuda.synchronize()
    assert int(length.item()) > cap                     # the caller can tell
    assert (text.cpu().numpy()[cap:] == 0x23).all()     # and nothing went past the capacity
    # beyond the built-in scratch: a clear error, nothing launched
    rc = lib.rt_format_ppm_noalloc(C.c_void_p(frame.data_ptr()), 7681, 4320, bkg, C.c_void_p(text.data_ptr()), cap,
                                   C.c_void_p(length.data_ptr()), None)
    assert rc == abi.RT_E_LIMIT
    assert b"33177600" in lib.rt_last_error()
