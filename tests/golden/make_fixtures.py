#!/usr/bin/env python
"""Generates the committed golden fixtures from the reference itself.

Run in the build container (needs /root/reference and oracle/_ref/, see
oracle/Makefile):

    python tests/golden/make_fixtures.py

Writes
  tests/golden/example_scenes.json   the text of the reference's example scene
                                      files (inputs; /root/reference does not
                                      exist on the GPU box)
  tests/golden/reference_frames.npz  what the UNMODIFIED reference program
                                      (oracle/_ref/raytracer_ref = g++ -O2
                                      -ffp-contract=off main.cpp) writes for each
                                      runnable scene, at a reduced `imsize` to keep
                                      the file small, plus the two full-size
                                      1080x1080 frames that the reference repo's
                                      own screenshots pin
  tests/golden/reference_frames.json sizes, NaN pixel counts, sha256 of the
                                      full-size frames, and the PNG check result

Scenes that need the absent Git-LFS textures get seeded stand-ins from
simple_raytracer_b200.scenegen written as P3 files next to the scene copy.
"""
from __future__ import annotations

import hashlib
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from simple_raytracer_b200 import scenegen  # noqa: E402

REFERENCE = "/root/reference"
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "raytracer_ref")
NAN_TOKEN = "18446744071562067968"

# (scene file, reduced imsize or None to keep, textures it needs)
SCENES = [
    ("basic_geometry_tests/four_spheres.txt", (128, 128), []),
    ("basic_geometry_tests/purple_pyramid.txt", (128, 64), []),
    ("lighting_tests/point_light.txt", (128, 128), []),
    ("lighting_tests/directional-light.txt", (128, 128), []),
    ("material_tests/beige_plastic.txt", (96, 96), []),
    ("material_tests/dull_copper.txt", (96, 96), []),
    ("material_tests/rubber_eraser.txt", (96, 96), []),
    ("material_tests/shiny_silver.txt", (96, 96), []),
    ("shadow_tests/shadow_test.txt", (64, 128), []),
    ("shadow_tests/multi-light-shadow.txt", (64, 64), []),
    ("reflection_transparency_tests/Test1.txt", (180, 180), []),
    ("reflection_transparency_tests/test7.txt", (180, 180), []),
    ("reflection_transparency_tests/Test2.txt", (120, 120), ["harbor.ppm"]),
    ("reflection_transparency_tests/Test3.txt", (120, 120), ["harbor.ppm"]),
    ("reflection_transparency_tests/Test4.txt", (120, 120), ["harbor.ppm"]),
    ("reflection_transparency_tests/Test5.txt", (120, 120), ["sunset.ppm"]),
    ("reflection_transparency_tests/Test6.txt", (120, 120), ["sunset.ppm"]),
    ("showcases/earth.txt", (128, 128), ["earthtexture.ppm"]),
    ("showcases/earth_pyramid.txt", (128, 128), ["earthtexture.ppm", "pyramid_texture.ppm"]),
    ("showcases/house.txt", (150, 100), ["grass.ppm", "wood.ppm", "redwood.ppm", "soccerball.ppm"]),
]
FULL_SIZE = {
    "reflection_transparency_tests/Test1.txt": "Examples/reflection_transparency_tests/Test1.png",
    "reflection_transparency_tests/test7.txt": "Examples/demo_image.png",
}


def set_imsize(text: str, size) -> str:
    if size is None:
        return text
    out = []
    for line in text.split("\n"):
        if line.startswith("imsize "):
            line = f"imsize {size[0]} {size[1]}"
        out.append(line)
    return "\n".join(out)


def read_ref_ppm(path: str):
    """-> raw int32 [H*W, 3] with INT32_MIN where the reference printed its NaN token"""
    with open(path) as f:
        tok = f.read().split()
    assert tok[0] == "P3" and tok[3] == "255", tok[:4]
    w, h = int(tok[1]), int(tok[2])
    body = tok[4:]
    assert len(body) == w * h * 3
    vals = np.array([-2147483648 if t == NAN_TOKEN else int(t) for t in body], dtype=np.int64)
    assert vals.min() >= -2147483648 and vals.max() <= 2147483647
    return w, h, vals.astype(np.int32).reshape(-1, 3)


def run_reference(scene_text: str, textures, workdir: str):
    os.makedirs(os.path.join(workdir, "textures"), exist_ok=True)
    for t in textures:
        # the scenes name their textures either "textures/x.ppm" or plain "x.ppm" (cwd-relative)
        for p in (os.path.join(workdir, "textures", t), os.path.join(workdir, t)):
            if not os.path.exists(p):
                with open(p, "w") as f:
                    f.write(scenegen.p3_text(scenegen.standin_texture(t)))
    with open(os.path.join(workdir, "scene.txt"), "w") as f:
        f.write(scene_text)
    r = subprocess.run([REF_BIN, "scene.txt"], cwd=workdir, capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stderr[-500:])
    return read_ref_ppm(os.path.join(workdir, "scene.ppm"))


def main() -> None:
    assert os.path.exists(REF_BIN), "build oracle/_ref first: make -C oracle"
    scenes, frames, meta = {}, {}, {}
    with tempfile.TemporaryDirectory() as tmp:
        for rel, size, textures in SCENES:
            with open(os.path.join(REFERENCE, "Examples", rel)) as f:
                text = f.read()
            key = rel[:-4]
            scenes[key] = {"text": text, "textures": textures}
            w, h, raw = run_reference(set_imsize(text, size), textures, tmp)
            frames[key] = raw.reshape(h, w, 3)
            meta[key] = {"width": w, "height": h, "nan_pixels": int((raw == -2147483648).any(axis=1).sum())}
            print(f"{key}: {w}x{h} nan_px={meta[key]['nan_pixels']}")
        try:
            from PIL import Image
        except ImportError:
            Image = None
        for rel, png in FULL_SIZE.items():
            key = rel[:-4]
            w, h, raw = run_reference(scenes[key]["text"], [], tmp)
            rgb = np.where(raw == -2147483648, 255, raw).astype(np.uint8)
            entry = {"width": w, "height": h, "nan_pixels": int((raw == -2147483648).any(axis=1).sum()),
                     "sha256_rgb_nan255": hashlib.sha256(rgb.tobytes()).hexdigest()}
            if Image is not None:
                im = np.asarray(Image.open(os.path.join(REFERENCE, png)).convert("RGB")).reshape(-1, 3)
                entry["png"] = png
                entry["png_max_abs_diff"] = int(np.abs(im.astype(np.int32) - rgb.astype(np.int32)).max())
                assert entry["png_max_abs_diff"] == 0, "reference screenshot no longer matches the reference build"
            meta[key + "@full"] = entry
            frames[key + "@full"] = raw.reshape(h, w, 3)
            print(f"{key}@full: {entry}")
    with open(os.path.join(HERE, "example_scenes.json"), "w") as f:
        json.dump(scenes, f, indent=1, sort_keys=True)
    # int16 is enough for 0..255 plus a NaN marker (-1); keeps the archive small
    packed = {k: np.where(v == -2147483648, -1, v).astype(np.int16) for k, v in frames.items()}
    np.savez_compressed(os.path.join(HERE, "reference_frames.npz"), **packed)
    with open(os.path.join(HERE, "reference_frames.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
