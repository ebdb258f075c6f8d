"""GPU parity at BASELINE.json's stated sizes, on the kernels and the launch schedule that
bench.py times: default parameters, NO instrumentation ("What's missing" 3).

  C1  four_spheres 512x512                          vs the CPU oracle
  C3  Test1, test7 at 1920x1080                     vs the CPU oracle
  C4  test7, house (stand-in textures) at 3840x2160 vs the CPU oracle
  C2  mesh_earth(512, 256) at 1920x1080             LBVH frame == brute-force frame (bytes)
  C5  1 M triangles + 1 k spheres at 3840x2160 and 7680x4320: tiles rendered by brute force
      (rank = k of world_size = 216 / 432, 10-row tiles) against the same rows of the full
      LBVH frame (two launch chains, merged any-hit launch, 8-wide tree), bytes; node counts
      per generation of the tile equal under both accelerators.

The oracle (oracle/rt_oracle.c) is pinned to the reference itself by tests/test_oracle_pinned.py.
"""
import numpy as np
import pytest

from simple_raytracer_b200 import abi, scenegen
from simple_raytracer_b200.renderer import DeviceScene
from tests import common
from tests import oracle_lib as O

pytestmark = pytest.mark.gpu

PIXEL_BAR = 0.999


def _plain_frame(scene, **kw):
    with DeviceScene(scene) as ds:
        st = ds.render(**kw)
        rgb, flags = ds.download()
    return rgb, flags, st


@pytest.mark.parametrize("name,w,h", [
    ("basic_geometry_tests/four_spheres", 512, 512),                    # C1
    ("lighting_tests/point_light", 512, 512),                           # C1, the "one point light" reading
    ("reflection_transparency_tests/Test1", 1920, 1080),                # C3
    ("reflection_transparency_tests/test7", 1920, 1080),                # C3
    ("reflection_transparency_tests/test7", 3840, 2160),                # C4 (i)
    ("showcases/house", 3840, 2160),                                    # C4 (ii), seeded stand-in textures
])
def test_example_scenes_at_baseline_sizes_match_the_oracle(name, w, h):
    scene = common.example_scene(name).with_size(w, h)
    raw, ids, ost = O.oracle_render(scene)
    rgb, flags, st = _plain_frame(scene)
    cmp = common.compare_frames(raw, rgb, flags)
    assert cmp["nan_mismatch"] == 0, cmp
    assert cmp["frac_within_1"] >= PIXEL_BAR and cmp["max_err"] <= 2, cmp
    # the wavefront's node / ray counts are the reference's ShadeRay / TraceRay call counts
    assert list(st.nodes_per_level)[:5] == list(ost.nodes_per_level)[:5]
    assert st.rays_transmit == ost.rays_transmit and st.rays_reflect == ost.rays_reflect
    assert st.rays_shadow == ost.rays_shadow and st.rays_primary == ost.rays_primary
    assert st.accel_used == abi.RT_ACCEL_BRUTE
    # ... and the LBVH pipeline on the same scene gives the same bytes
    rgb2, flags2, st2 = _plain_frame(scene, accel=abi.RT_ACCEL_LBVH)
    assert np.array_equal(rgb2, rgb) and np.array_equal(flags2, flags)
    assert list(st2.nodes_per_level) == list(st.nodes_per_level)


def test_textured_mesh_at_1080p_lbvh_equals_brute_force():
    """C2 at its full size: 261,120 smooth textured triangles, directional + point light."""
    scene = scenegen.mesh_earth(512, 256, 1920, 1080)
    assert len(scene.faces) == 261_120
    with DeviceScene(scene) as ds:
        st = ds.render()
        assert st.accel_used == abi.RT_ACCEL_LBVH
        rgb, flags = ds.download()
        rgb, flags = rgb.copy(), flags.copy()
        # brute force on every 9th tile of 8 rows (each tile costs 15,360 pixels x 261 k primitives x 3 rays)
        for rank in (0, 4, 8):
            got_rgb = np.zeros_like(rgb); got_flags = np.full_like(flags, 255)
            sb = ds.render(accel=abi.RT_ACCEL_BRUTE, rank=rank, world_size=9, tile_rows=8)
            ds.download(got_rgb, got_flags)
            sl = ds.render(accel=abi.RT_ACCEL_LBVH, rank=rank, world_size=9, tile_rows=8)
            rows = np.array([r for r in range(1080) if (r // 8) % 9 == rank])
            assert np.array_equal(got_rgb[rows], rgb[rows]) and np.array_equal(got_flags[rows], flags[rows]), rank
            assert list(sb.nodes_per_level) == list(sl.nodes_per_level)
    # the CPU oracle on a band of rows of the same frame (the whole frame would take minutes)
    pix = (np.arange(530, 534)[:, None] * 1920 + np.arange(1920)[None, :]).reshape(-1)
    raw, ids, _ = O.oracle_render(scene, pixels=pix)
    cmp = common.compare_frames(raw, rgb.reshape(-1, 3)[pix], flags.reshape(-1)[pix])
    assert cmp["nan_mismatch"] == 0 and cmp["frac_within_1"] >= PIXEL_BAR and cmp["max_err"] <= 2, cmp


@pytest.mark.parametrize("w,h,world,tiles", [(3840, 2160, 216, (0, 107, 215)), (7680, 4320, 432, (3, 216, 431))])
def test_stress_frame_tiles_by_brute_force(w, h, world, tiles):
    """C5: the production LBVH frame against brute-force renders of three of its 10-row tiles."""
    scene = scenegen.stress_scene(1_000_000, 1000, w, h)
    with DeviceScene(scene) as ds:
        st = ds.render(tile_rows=10)                       # default schedule: K = 2 chains at these sizes
        assert st.accel_used == abi.RT_ACCEL_LBVH and st.queue_regrows >= 0
        assert st.bvh_levels4 > 0 and 3 * st.bvh_levels4 + 2 <= 96      # the static stack bound of build_bvh
        rgb, flags = ds.download()
        rgb, flags = rgb.copy(), flags.copy()
        for k in tiles:
            got_rgb = np.zeros_like(rgb); got_flags = np.full_like(flags, 255)
            sb = ds.render(accel=abi.RT_ACCEL_BRUTE, rank=k, world_size=world, tile_rows=10)
            ds.download(got_rgb, got_flags)
            sl = ds.render(accel=abi.RT_ACCEL_LBVH, rank=k, world_size=world, tile_rows=10)
            rows = np.arange(k * 10, k * 10 + 10)
            assert np.array_equal(got_rgb[rows], rgb[rows]), (k, int((got_rgb[rows] != rgb[rows]).any(axis=2).sum()))
            assert np.array_equal(got_flags[rows], flags[rows]), k
            assert list(sb.nodes_per_level) == list(sl.nodes_per_level), k
            assert sb.rays_transmit == sl.rays_transmit and sb.rays_reflect == sl.rays_reflect
        # instrumented frame, same schedule: same bytes, and the ray counts the bench quotes
        si = ds.render(tile_rows=10, instrument=True)
        again = ds.download()
        assert np.array_equal(again[0], rgb) and np.array_equal(again[1], flags)
        assert list(si.nodes_per_level) == list(st.nodes_per_level)
        assert 0 < si.shadow_rays_walked <= si.rays_shadow
        # the occluder maps (bvh.cu: build_shadow_maps) settle most of this dense scene's shadow rays without a walk
        assert si.shadow_map_hits > 3 * si.shadow_rays_walked, (si.shadow_map_hits, si.shadow_rays_walked)


def test_shadow_probe_on_the_8_wide_tree_is_bit_exact():
    """The any-hit kernel's own tree (collapse8_kernel, node-ordered primitive records) through the
    probe: shadow products and occluder counts against the oracle and, on the 1 M scene, brute force."""
    scene = scenegen.random_soup(400, 40, 9, 32, 24)
    rng = np.random.default_rng(1)
    n = 20000
    o = rng.uniform(-12, 12, (n, 3)).astype(np.float32)
    d = rng.normal(0, 1, (n, 3)).astype(np.float32)
    d[: n // 2] /= np.linalg.norm(d[: n // 2], axis=1, keepdims=True)
    self_id = rng.integers(1, scene.n_prims + 1, n).astype(np.int32)
    tmax = rng.uniform(1, 30, n).astype(np.float32)
    with DeviceScene(scene) as ds:
        for tm in (None, tmax):
            m, h = O.oracle_trace_shadow(scene, o, d, self_id, 1.0e-3, tm)
            gm, gh = ds.trace_shadow(o, d, self_id, 1.0e-3, tm, accel=abi.RT_ACCEL_LBVH8)
            assert (h > 0).sum() > 1000
            assert np.array_equal(gh, h) and np.array_equal(gm.view(np.uint32), m.view(np.uint32))
    big = scenegen.stress_scene(1_000_000, 1000, 64, 48)
    rng = np.random.default_rng(12)
    n = 2048
    o = rng.uniform(-70, 70, (n, 3)).astype(np.float32)
    d = rng.normal(0, 1, (n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    self_id = rng.integers(1, big.n_prims + 1, n).astype(np.int32)
    with DeviceScene(big) as ds:
        a = ds.trace_shadow(o, d, self_id, 1.0e-3, None, accel=abi.RT_ACCEL_BRUTE)
        b = ds.trace_shadow(o, d, self_id, 1.0e-3, None, accel=abi.RT_ACCEL_LBVH8)
    assert (a[1] > 0).sum() > 500
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32))


def _panes_scene(n_panes, width=96, height=72, opacity=0.05):
    """A floor under a stack of `n_panes` slightly opaque quads, lit from (almost) straight above by a
    directional light and from the side by a point light: a shadow ray from the floor under the
    stack meets n_panes partial occluders (a pane is two triangles; the ray crosses one of them),
    whose factors (float)(1 - 0.05) multiply to a value that depends on the order of the products."""
    from simple_raytracer_b200.scene import Scene
    mats = np.array([scenegen.material((0.9, 0.9, 0.9), (1, 1, 1), 0.2, 0.7, 0.1, 10),
                     scenegen.material((0.3, 0.6, 0.9), (1, 1, 1), 0.1, 0.3, 0.0, 20, opacity, 1.0)],
                    dtype=scenegen.MATERIAL_DTYPE)
    faces = []

    def quad(p0, p1, p2, p3, mat):
        for a, b, c in ((0, 1, 2), (0, 2, 3)):
            f = np.zeros(1, dtype=scenegen.FACE_DTYPE)[0]
            P = [p0, p1, p2, p3]
            f["v"] = [P[a], P[b], P[c]]
            f["vn"] = [(0, 1, 0)] * 3
            f["smooth"], f["material"], f["texture"] = 0, mat, -1
            faces.append(f)

    quad((-6, 0, 4), (6, 0, 4), (6, 0, 16), (-6, 0, 16), 0)
    for k in range(n_panes):
        # shuffled file order: the product must follow iteration order, not height
        y = 1.0 + 0.05 * ((k * 7) % n_panes)
        quad((-2, y, 8), (2, y, 8), (2, y, 12), (-2, y, 12), 1)
    faces = np.array(faces, dtype=scenegen.FACE_DTYPE)
    faces["id"] = np.arange(1, len(faces) + 1)
    lights = np.array([scenegen.light((0.05, -1, 0.02), 0, (0.8, 0.8, 0.8)), scenegen.light((8, 6, 3), 1, (0.4, 0.4, 0.4))],
                      dtype=scenegen.LIGHT_DTYPE)
    return Scene(width, height, eye=(0, 5, -3), viewdir=(0, -0.45, 1), updir=(0, 1, 0), hfov_deg=60, bkg=(0.2, 0.3, 0.5),
                 bkg_eta=1.0, materials=mats, faces=faces, lights=lights)


@pytest.mark.parametrize("n_panes", [10, 30, 150])
def test_many_stacked_transparent_occluders_keep_the_reference_product(n_panes):
    """More partial occluders on one shadow ray than the any-hit kernel's visitor holds (12): the
    node's chain goes through bvh_shadow_redo_kernel, whose product follows the reference's
    iteration order for ANY count — frames equal the brute-force frame and the oracle's."""
    scene = _panes_scene(n_panes)
    raw, ids, ost = O.oracle_render(scene, max_depth=1)
    with DeviceScene(scene) as ds:
        ds.render(accel=abi.RT_ACCEL_BRUTE, max_depth=1)
        want = ds.download()
        want = (want[0].copy(), want[1].copy())
        for sharing in (2, 1):
            st = ds.render(accel=abi.RT_ACCEL_LBVH, max_depth=1, work_sharing=sharing)
            got = ds.download()
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (n_panes, sharing)
            if n_panes > 12:
                assert st.shadow_chains_redone > 0
            else:
                assert st.shadow_chains_redone == 0
    cmp = common.compare_frames(raw, want[0], want[1])
    assert cmp["nan_mismatch"] == 0 and cmp["frac_within_1"] >= PIXEL_BAR and cmp["max_err"] <= 2, cmp


def test_coincident_primitives_do_not_break_the_tree():
    """2,000 copies of one triangle and 300 copies of one sphere among a small soup: equal Morton
    codes down to the last bit (the LBVH splits them by index), equal t for every copy (the tie rule
    picks the first in file order).  LBVH frame == brute-force frame; the tree stays shallow enough
    for the walk stacks (build_bvh would refuse it otherwise)."""
    base = scenegen.random_soup(300, 10, 33, 96, 72)
    tri = np.repeat(base.faces[:1], 2000)
    sph = np.repeat(base.spheres[:1], 300)
    faces = np.concatenate([base.faces, tri]); spheres = np.concatenate([base.spheres, sph])
    spheres["id"] = np.arange(1, len(spheres) + 1)
    faces["id"] = np.arange(len(spheres) + 1, len(spheres) + len(faces) + 1)
    scene = base.copy(faces=faces, spheres=spheres)
    with DeviceScene(scene) as ds:
        sb = ds.render(accel=abi.RT_ACCEL_BRUTE)
        want = ds.download()
        want = (want[0].copy(), want[1].copy())
        sl = ds.render(accel=abi.RT_ACCEL_LBVH, instrument=True)
        got = ds.download()
        ids_l = ds.download_primary_ids().copy()
        ds.render(accel=abi.RT_ACCEL_BRUTE, instrument=True)
        ids_b = ds.download_primary_ids()
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    assert np.array_equal(ids_l, ids_b)
    assert list(sb.nodes_per_level) == list(sl.nodes_per_level)
    assert 0 < sl.bvh_levels4 <= 31


@pytest.mark.parametrize("name", ["reflection_transparency_tests/Test1", "showcases/house", "showcases/earth_pyramid",
                                  "basic_geometry_tests/purple_pyramid"])
def test_reference_program_with_the_drop_in_binding_writes_the_reference_file(tmp_path, name):
    """oracle/_ref/raytracer_b200_dropin = the reference's own main.cpp (parser, texture loader, PPM
    writer) with the call at main.cpp:607 served by librt_b200.so through the binding of INTEGRATION.md
    (oracle/ref_patched.cpp), next to the unmodified reference program on the same scene file and
    textures: same header and line structure, the same NaN tokens in the same places, every channel
    within 1 LSB on >= 99.9 % of the pixels (measured: all of them equal)."""
    import os, subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    dropin = os.path.join(root, "oracle", "_ref", "raytracer_b200_dropin")
    ref = os.path.join(root, "oracle", "_ref", "raytracer_ref")
    if not (os.path.exists(dropin) and os.path.exists(ref)):
        pytest.skip("oracle/_ref/raytracer_b200_dropin or raytracer_ref not built")
    scene, _ = common.golden_frame(name)
    outs = {}
    for who, exe in (("theirs", ref), ("ours", dropin)):
        d = tmp_path / who
        d.mkdir()
        path = common.materialise(name, str(d))
        text = open(path).read()
        text = "\n".join(f"imsize {scene.width} {scene.height}" if l.startswith("imsize ") else l for l in text.split("\n"))
        (d / "scene.txt").write_text(text)
        subprocess.run([exe, "scene.txt"], cwd=str(d), check=True, capture_output=True, timeout=600)
        outs[who] = (d / "scene.ppm").read_bytes().split(b"\n")
    a, b = outs["ours"], outs["theirs"]
    assert a[:3] == b[:3] and len(a) == len(b) and a[-1] == b[-1] == b""
    nan = b"18446744071562067968"
    va = np.array([[-1 if t == nan else int(t) for t in l.split()] for l in a[3:-1]], dtype=np.int64)
    vb = np.array([[-1 if t == nan else int(t) for t in l.split()] for l in b[3:-1]], dtype=np.int64)
    assert va.shape == vb.shape == (scene.width * scene.height, 3)
    assert np.array_equal(va == -1, vb == -1)
    ok = (np.abs(va - vb) <= 1).all(axis=1)
    assert ok.mean() >= PIXEL_BAR, float(ok.mean())
    assert int(np.abs(va - vb).max()) <= 2
