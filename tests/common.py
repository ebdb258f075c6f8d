"""Shared test helpers: example scenes and golden frames from tests/golden/."""
from __future__ import annotations

import functools
import json
import os
import tempfile

import numpy as np

from simple_raytracer_b200 import scenegen
from simple_raytracer_b200.scene import Scene

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")
NAN = -2147483648

TEXTURE_FREE = [
    "basic_geometry_tests/four_spheres", "basic_geometry_tests/purple_pyramid", "lighting_tests/point_light",
    "lighting_tests/directional-light", "material_tests/beige_plastic", "material_tests/dull_copper",
    "material_tests/rubber_eraser", "material_tests/shiny_silver", "shadow_tests/shadow_test",
    "shadow_tests/multi-light-shadow", "reflection_transparency_tests/Test1", "reflection_transparency_tests/test7",
]
TEXTURED = [
    "reflection_transparency_tests/Test2", "reflection_transparency_tests/Test3", "reflection_transparency_tests/Test4",
    "reflection_transparency_tests/Test5", "reflection_transparency_tests/Test6", "showcases/earth",
    "showcases/earth_pyramid", "showcases/house",
]
ALL_SCENES = TEXTURE_FREE + TEXTURED


@functools.lru_cache(maxsize=None)
def _scene_table() -> dict:
    with open(os.path.join(GOLDEN, "example_scenes.json")) as f:
        return json.load(f)


@functools.lru_cache(maxsize=None)
def _frame_meta() -> dict:
    with open(os.path.join(GOLDEN, "reference_frames.json")) as f:
        return json.load(f)


@functools.lru_cache(maxsize=None)
def _frames():
    return np.load(os.path.join(GOLDEN, "reference_frames.npz"))


def scene_text(name: str) -> str:
    return _scene_table()[name]["text"]


def materialise(name: str, workdir: str, tex_scale: int = 1) -> str:
    """Writes the scene file and the stand-in textures it names; returns the scene path."""
    entry = _scene_table()[name]
    os.makedirs(os.path.join(workdir, "textures"), exist_ok=True)
    for t in entry["textures"]:
        for p in (os.path.join(workdir, "textures", t), os.path.join(workdir, t)):
            if not os.path.exists(p):
                with open(p, "w") as f:
                    f.write(scenegen.p3_text(scenegen.standin_texture(t, tex_scale)))
    path = os.path.join(workdir, name.replace("/", "__") + ".txt")
    with open(path, "w") as f:
        f.write(entry["text"])
    return path


@functools.lru_cache(maxsize=None)
def example_scene(name: str, tex_scale: int = 1) -> Scene:
    """The example scene parsed by OUR host front-end at its own imsize (tex_scale: the seeded
    stand-in textures that many times larger than the 128-wide ones the fixtures were made with)."""
    with tempfile.TemporaryDirectory() as tmp:
        return Scene.from_file(materialise(name, tmp, tex_scale), texture_dir=tmp)


def golden_frame(name: str):
    """-> (Scene at the fixture's imsize, raw int32 [H*W,3] of the UNMODIFIED reference program)"""
    meta = _frame_meta()[name]
    base = name[:-5] if name.endswith("@full") else name
    sc = example_scene(base).with_size(meta["width"], meta["height"])
    raw = _frames()[name].astype(np.int32).reshape(-1, 3)
    raw = np.where(raw == -1, NAN, raw)
    return sc, raw


def frame_meta(name: str) -> dict:
    return _frame_meta()[name]


def compare_frames(raw_ref: np.ndarray, rgb: np.ndarray, flags: np.ndarray) -> dict:
    """North-star pixel bar: +-1 LSB per channel on >= 99.9 % of pixels, max error reported,
    NaN pixel set identical.  raw_ref: int32 [n,3] (NAN marker); rgb uint8 [n,3]; flags uint8 [n]."""
    rgb = rgb.reshape(-1, 3).astype(np.int32)
    flags = flags.reshape(-1)
    ref_nan = raw_ref == NAN
    got_nan = np.stack([(flags & 1) != 0, (flags & 2) != 0, (flags & 4) != 0], axis=1)
    both = ~ref_nan & ~got_nan
    diff = np.where(both, np.abs(np.clip(raw_ref, 0, 255) - rgb), 0)
    px_err = diff.max(axis=1)
    nan_mismatch = (ref_nan != got_nan).any(axis=1)
    n = len(raw_ref)
    return {
        "pixels": n,
        "exact": int(((px_err == 0) & ~nan_mismatch).sum()),
        "within_1": int(((px_err <= 1) & ~nan_mismatch).sum()),
        "frac_within_1": float(((px_err <= 1) & ~nan_mismatch).sum() / n),
        "max_err": int(px_err.max()) if n else 0,
        "nan_ref": int(ref_nan.any(axis=1).sum()),
        "nan_mismatch": int(nan_mismatch.sum()),
    }
