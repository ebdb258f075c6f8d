"""N GPUs behind the C ABI (rt_upload_scene_multi / rt_render_multi, include/rt.h): ONE process drives
all devices, NCCL lives inside librt_b200.so.  The frame must be byte-identical for every N.
N = 1 exercises the whole code path except the collectives and runs on any GPU box; N >= 2 needs that
many devices (the driver's single-GPU box skips those; tools/r2_multi.sh runs them on 2 / 8 GPUs)."""
import os
import subprocess

import numpy as np
import pytest

from simple_raytracer_b200 import abi, scenegen
from simple_raytracer_b200.renderer import DeviceScene, MultiDeviceScene
from tests import common

pytestmark = pytest.mark.gpu


def _n_devices() -> int:
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _single(scene, **kw):
    with DeviceScene(scene) as ds:
        st = ds.render(**kw)
        rgb, flags = ds.download()
        text = ds.ppm_text()
    return rgb.copy(), flags.copy(), st, text


@pytest.mark.parametrize("n", [1, 2, 4, 8])
@pytest.mark.parametrize("make", [lambda: scenegen.coverage_scene(160, 101),              # brute force, ragged tiles
                                  lambda: scenegen.stress_scene(20000, 100, 320, 200),     # LBVH built once, broadcast
                                  lambda: common.golden_frame("reflection_transparency_tests/Test1")[0]])   # NaN pixels
def test_one_process_n_gpus_give_the_single_gpu_frame(make, n):
    if _n_devices() < n:
        pytest.skip(f"needs {n} CUDA devices")
    scene = make()
    rgb, flags, st1, text1 = _single(scene)
    with MultiDeviceScene(scene, n) as ms:
        assert ms.num_gpus == n
        for k in range(4):                       # direct, capture, replay, replay on every device
            st = ms.render()
            got = ms.download()
            assert np.array_equal(got[0], rgb) and np.array_equal(got[1], flags), (n, k)
            assert list(st.nodes_per_level) == list(st1.nodes_per_level)
            assert st.rays_primary == st1.rays_primary and st.rays_reflect == st1.rays_reflect
        assert st.graph_replays >= n
        assert ms.ppm_text() == text1
        # explicit tile height, odd on purpose
        ms.render(tile_rows=7)
        got = ms.download()
        assert np.array_equal(got[0], rgb) and np.array_equal(got[1], flags)


@pytest.mark.parametrize("n", [1, 2, 8])
def test_cli_gpus_n_writes_the_same_file(tmp_path, n):
    """`bin/raytracer --gpus N scene.txt` (the C++ host; no Python, no torch in that process)."""
    if _n_devices() < n:
        pytest.skip(f"needs {n} CUDA devices")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cli = os.path.join(root, "bin", "raytracer")
    if not os.path.exists(cli):
        pytest.skip("bin/raytracer not built")
    outs = []
    for k, g in enumerate((1, n)):
        d = tmp_path / f"run{k}_g{g}"
        d.mkdir()
        path = common.materialise("showcases/house", str(d))
        os.rename(path, str(d / "scene.txt"))
        r = subprocess.run([cli, "--gpus", str(g), "--frames", "3", "--stats", "scene.txt"], cwd=str(d), capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr
        assert f"gpus {g} " in r.stderr
        outs.append((d / "scene.ppm").read_bytes())
    assert outs[0] == outs[1] and len(outs[0]) > 1000


def test_more_gpus_than_visible_is_an_error():
    scene = scenegen.coverage_scene(32, 24)
    with pytest.raises(abi.RtError) as e:
        MultiDeviceScene(scene, _n_devices() + 1)
    assert e.value.code == abi.RT_E_INVALID
