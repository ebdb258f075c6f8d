"""Pins the CPU oracle (oracle/rt_oracle.c) to the reference.

  1. committed golden frames written by the UNMODIFIED reference program
     (tests/golden/reference_frames.npz, made by tests/golden/make_fixtures.py)
     for every runnable example scene, reached through OUR parser: bit-exact;
  2. the two full-size frames that the reference repository's own screenshots
     pin (Test1.png, demo_image.png): bit-exact, checked through a sha256 of the
     frame with NaN pixels shown as 255;
  3. when oracle/_ref/libref.so is present (build container; it also travels to
     the GPU box): bit-exact against the reference's own ShadeRay / TraceRay on
     generated scenes covering transparent faces, textures, nested media.
"""
import hashlib

import numpy as np
import pytest

from simple_raytracer_b200 import scenegen
from tests import common
from tests import oracle_lib as O


@pytest.mark.parametrize("name", common.ALL_SCENES)
def test_oracle_matches_reference_program(name):
    scene, ref_raw = common.golden_frame(name)
    raw, ids, st = O.oracle_render(scene)
    assert raw.shape == ref_raw.shape
    bad = int((raw != ref_raw).any(axis=1).sum())
    assert bad == 0, f"{name}: {bad} pixels differ from the reference program"
    assert int((raw == O.RTO_NAN).any(axis=1).sum()) == common.frame_meta(name)["nan_pixels"]


@pytest.mark.parametrize("name", ["reflection_transparency_tests/Test1", "reflection_transparency_tests/test7"])
def test_oracle_matches_reference_screenshots(name):
    meta = common.frame_meta(name + "@full")
    assert meta.get("png_max_abs_diff", 0) == 0          # the PNG check done when the fixture was made
    scene = common.example_scene(name)
    assert (scene.width, scene.height) == (meta["width"], meta["height"])
    raw, _, _ = O.oracle_render(scene)
    assert int((raw == O.RTO_NAN).any(axis=1).sum()) == meta["nan_pixels"]
    rgb = np.where(raw == O.RTO_NAN, 255, raw).astype(np.uint8)
    assert hashlib.sha256(rgb.tobytes()).hexdigest() == meta["sha256_rgb_nan255"]
    # and the stored full frame itself
    _, ref_raw = common.golden_frame(name + "@full")
    assert np.array_equal(raw, ref_raw)


def test_ray_counts_follow_the_reference_unit():
    """One ray = one TraceRay call: primary = pixels, shadow = nodes * lights."""
    scene, _ = common.golden_frame("reflection_transparency_tests/Test1")
    _, _, st = O.oracle_render(scene)
    assert st.rays_primary == scene.n_pixels
    assert st.rays_shadow == st.nodes_total * len(scene.lights)
    assert st.nodes_total == sum(st.nodes_per_level)
    assert st.rays_transmit + st.rays_reflect >= st.nodes_total - st.nodes_per_level[0]


def test_depth_zero_has_no_secondary_rays():
    scene, _ = common.golden_frame("lighting_tests/point_light")
    _, _, st = O.oracle_render(scene, max_depth=0)
    assert st.rays_transmit == 0 and st.rays_reflect == 0
    assert st.nodes_total == st.nodes_per_level[0]


needs_ref = pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref/libref.so not built")


@needs_ref
@pytest.mark.parametrize("make", [
    lambda: scenegen.coverage_scene(96, 72),
    lambda: scenegen.random_soup(300, 20, 1, 64, 48),
    lambda: scenegen.random_soup(60, 6, 5, 64, 48),
    lambda: scenegen.mesh_earth(16, 8, 64, 36, tex_scale=1),
    lambda: common.example_scene("showcases/house").with_size(90, 60),
])
def test_oracle_matches_reference_code_on_generated_scenes(make):
    scene = make()
    raw, ids, _ = O.oracle_render(scene)
    ref_raw, _ = O.ref_render(scene)
    assert np.array_equal(raw, ref_raw)
    # the shim's per-pixel loop (used for timing subsets) equals the reference's own frame loop
    px = np.arange(0, scene.n_pixels, 7, dtype=np.int64)
    sub_raw, sub_ids, _ = O.ref_render_pixels(scene, px)
    assert np.array_equal(sub_raw, ref_raw[px])
    assert np.array_equal(sub_ids, ids[px])


@needs_ref
@pytest.mark.parametrize("depth", [0, 1, 2, 3])
def test_oracle_matches_reference_code_at_other_depths(depth):
    scene, _ = common.golden_frame("reflection_transparency_tests/test7")
    raw, _, _ = O.oracle_render(scene, max_depth=depth)
    ref_raw, _ = O.ref_render(scene, max_depth=depth)
    assert np.array_equal(raw, ref_raw)


@needs_ref
def test_trace_probes_match_reference_traceray():
    """Hit ids, t bits, normals, barycentrics and shadow products of random rays."""
    scene = scenegen.random_soup(200, 30, 9, 32, 24)
    rng = np.random.default_rng(0)
    n = 4000
    o = rng.uniform(-12, 12, (n, 3)).astype(np.float32)
    d = rng.normal(0, 1, (n, 3)).astype(np.float32)
    d[: n // 2] /= np.linalg.norm(d[: n // 2], axis=1, keepdims=True)      # half unit, half not (A3.2)
    ids, t, nrm, bary = O.oracle_trace_closest(scene, o, d, 1.0e-3)
    rids, rt_, rnrm, rbary = O.ref_trace_closest(scene, o, d, 1.0e-3)
    assert np.array_equal(ids, rids) and (ids > 0).sum() > 100
    assert np.array_equal(t.view(np.uint32), rt_.view(np.uint32))
    hit = ids > 0
    assert np.array_equal(nrm[hit].view(np.uint32), rnrm[hit].view(np.uint32))
    assert np.array_equal(bary[hit].view(np.uint32), rbary[hit].view(np.uint32))
    self_id = rng.integers(1, scene.n_prims + 1, n).astype(np.int32)
    tmax = rng.uniform(1, 30, n).astype(np.float32)
    for tm in (None, tmax):
        m, h = O.oracle_trace_shadow(scene, o, d, self_id, 1.0e-3, tm)
        rm, rh = O.ref_trace_shadow(scene, o, d, self_id, 1.0e-3, tm)
        assert np.array_equal(h, rh) and np.array_equal(m.view(np.uint32), rm.view(np.uint32))
        assert (h > 0).sum() > 100
