"""Seeded synthetic inputs: stand-in textures for the Git-LFS files the reference
repository does not carry, and the generated scenes of
BASELINE.json's configs 2 and 5.  Everything is a pure function of
its seed so the build container and the GPU box generate identical bytes.
"""
from __future__ import annotations

import numpy as np

from .scene import FACE_DTYPE, LIGHT_DTYPE, MATERIAL_DTYPE, SPHERE_DTYPE, Scene


# ---- textures -------------------------------------------------------------------
def value_noise(h: int, w: int, seed: int, cells: int = 8) -> np.ndarray:
    """Smooth noise in [0,1], periodic in x (so lat-long textures have no seam)."""
    rng = np.random.default_rng(seed)
    out = np.zeros((h, w), dtype=np.float64)
    amp, total = 1.0, 0.0
    for octave in range(4):
        c = cells << octave
        g = rng.random((c + 1, c))
        g = np.concatenate([g, g[:, :1]], axis=1)          # periodic in x
        ys = np.linspace(0, c, h, endpoint=False)
        xs = np.linspace(0, c, w, endpoint=False)
        y0, x0 = ys.astype(int), xs.astype(int)
        fy, fx = ys - y0, xs - x0
        fy, fx = fy * fy * (3 - 2 * fy), fx * fx * (3 - 2 * fx)
        a = g[np.ix_(y0, x0)]; b = g[np.ix_(y0, x0 + 1)]
        cc = g[np.ix_(y0 + 1, x0)]; d = g[np.ix_(y0 + 1, x0 + 1)]
        out += amp * ((a * (1 - fx) + b * fx) * (1 - fy)[:, None] + (cc * (1 - fx) + d * fx) * fy[:, None])
        total += amp
        amp *= 0.5
    return out / total


def make_texture(kind: str, width: int, height: int, seed: int) -> np.ndarray:
    """uint8 [height][width][3].  Hard thresholds and stripes on purpose: a wrong
    texel index shows up as a large pixel error instead of hiding in smooth noise."""
    n = value_noise(height, width, seed)
    yy, xx = np.mgrid[0:height, 0:width]
    if kind == "earth":
        land = n > 0.5
        r = np.where(land, 60 + 120 * n, 10 + 30 * n)
        g = np.where(land, 120 + 100 * n, 40 + 80 * n)
        b = np.where(land, 40 + 60 * n, 120 + 130 * n)
    elif kind == "grass":
        r, g, b = 30 + 60 * n, 110 + 120 * n, 20 + 50 * n
    elif kind == "wood":
        s = 0.5 + 0.5 * np.sin(xx * 0.2 + 6 * n)
        r, g, b = 120 + 80 * s, 70 + 60 * s, 30 + 30 * s
    elif kind == "redwood":
        s = ((yy // max(1, height // 16)) % 2).astype(np.float64)
        r, g, b = 140 + 80 * s + 20 * n, 40 + 20 * s, 30 + 20 * n
    elif kind == "soccerball":
        s = (((xx // max(1, width // 8)) + (yy // max(1, height // 8))) % 2).astype(np.float64)
        r = g = b = 20 + 220 * s
    elif kind == "checker":
        s = (((xx // max(1, width // 16)) + (yy // max(1, height // 16))) % 2).astype(np.float64)
        r, g, b = 40 + 200 * s, 200 - 150 * s, 60 + 120 * n
    else:   # sky-like panoramas: harbor, sunset, ...
        t = yy / max(1, height - 1)
        r, g, b = 240 - 120 * t + 30 * n, 150 + 60 * t * n, 90 + 140 * t
    img = np.stack([r, g, b], axis=-1)
    return np.clip(np.round(img), 0, 255).astype(np.uint8)


# file name (as the example scenes spell it) -> (kind, width, height, seed)
STANDIN_TEXTURES = {
    "earthtexture.ppm": ("earth", 512, 256, 2024),
    "pyramid_texture.ppm": ("checker", 128, 128, 7),
    "grass.ppm": ("grass", 128, 128, 1),
    "wood.ppm": ("wood", 128, 128, 2),
    "redwood.ppm": ("redwood", 128, 128, 3),
    "soccerball.ppm": ("soccerball", 128, 64, 4),
    "harbor.ppm": ("harbor", 256, 128, 5),
    "sunset.ppm": ("sunset", 256, 128, 6),
}


def standin_texture(file_name: str, scale: int = 1) -> np.ndarray:
    kind, w, h, seed = STANDIN_TEXTURES[file_name]
    return make_texture(kind, w * scale, h * scale, seed)


def p3_text(img: np.ndarray) -> str:
    """P3 text the reference loader accepts (utility.h:59-139): no empty lines,
    single spaces, one image row per line."""
    h, w, _ = img.shape
    rows = [" ".join(map(str, img[y].reshape(-1))) for y in range(h)]
    return f"P3\n{w} {h}\n255\n" + "\n".join(rows) + "\n"


# ---- helpers ------------------------------------------------------------------------
def material(od, os_, ka, kd, ks, n, opacity=1.0, eta=1.0):
    m = np.zeros(1, dtype=MATERIAL_DTYPE)[0]
    m["od"], m["os"] = od, os_
    m["ka"], m["kd"], m["ks"], m["n"], m["opacity"], m["eta"] = ka, kd, ks, n, opacity, eta
    return m


def light(xyz, w, rgb):
    l = np.zeros(1, dtype=LIGHT_DTYPE)[0]
    l["xyz"], l["w"], l["rgb"] = xyz, w, rgb
    return l


# ---- config 2: smooth-shaded textured lat-long mesh ------------------------------------
def mesh_earth(nlon: int = 64, nlat: int = 32, width: int = 1920, height: int = 1080,
               tex_scale: int = 4) -> Scene:
    """`f v/vt/vn` sphere of radius 2 at the origin, camera and material of
    Examples/showcases/earth.txt, one directional + one point light.
    2*nlon*(nlat-1) triangles: 64x32 -> 3,968; 512x256 -> 261,120."""
    lon = np.linspace(0.0, 2.0 * np.pi, nlon + 1)
    lat = np.linspace(0.0, np.pi, nlat + 1)
    LON, LAT = np.meshgrid(lon, lat)                         # [nlat+1][nlon+1]
    nrm = np.stack([np.sin(LAT) * np.cos(LON), np.sin(LAT) * np.sin(LON), np.cos(LAT)], axis=-1)
    pos = (2.0 * nrm).astype(np.float32)
    nrm = nrm.astype(np.float32)
    uv = np.stack([LON / (2.0 * np.pi), LAT / np.pi], axis=-1).astype(np.float32)
    i, j = np.meshgrid(np.arange(nlat), np.arange(nlon), indexing="ij")
    a = (i, j); b = (i + 1, j); c = (i + 1, j + 1); d = (i, j + 1)
    tris = []
    for (p, q, r), keep in (((a, b, c), i < nlat - 1), ((a, c, d), i > 0)):   # drop the degenerate pole triangles
        k = keep.reshape(-1)
        tris.append(tuple(np.stack([x[t].reshape(-1, x.shape[-1])[k] for t in (p, q, r)], axis=1)
                          for x in (pos, nrm, uv)))
    V = np.concatenate([t[0] for t in tris]); N = np.concatenate([t[1] for t in tris]); T = np.concatenate([t[2] for t in tris])
    faces = np.zeros(len(V), dtype=FACE_DTYPE)
    faces["v"], faces["vn"], faces["vt"] = V, N, T
    faces["smooth"], faces["material"], faces["texture"] = 1, 0, 0
    faces["id"] = np.arange(1, len(V) + 1)
    mats = np.array([material((0, 1, 0), (1, 1, 1), 0.2, 0.8, 0.1, 20)], dtype=MATERIAL_DTYPE)
    lights = np.array([light((0, 1, -1), 0, (1, 1, 1)), light((5, -6, 4), 1, (0.8, 0.8, 0.8))], dtype=LIGHT_DTYPE)
    return Scene(width, height, eye=(5, -5, 3), viewdir=(-5, 5, -3), updir=(0, 0, 1), hfov_deg=45,
                 bkg=(0.1, 0.1, 0.1), bkg_eta=0.0, materials=mats, faces=faces, lights=lights,
                 textures=[standin_texture("earthtexture.ppm", tex_scale)])


# ---- config 5: synthetic stress scene ---------------------------------------------------
def stress_scene(n_tris: int = 1_000_000, n_spheres: int = 1000, width: int = 7680, height: int = 4320,
                 seed: int = 12345, n_lights: int = 8, tex_size: int = 1024) -> Scene:
    """Random textured triangles + spheres, 4 directional (normalised) + 4 point
    lights, 16 materials of which 4 are transparent, 4 textures."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    c = rng.uniform(-50, 50, (n_tris, 1, 3))
    V = (c + rng.uniform(-0.6, 0.6, (n_tris, 3, 3))).astype(f32)
    e1 = V[:, 1] - V[:, 0]; e2 = V[:, 2] - V[:, 0]
    fn = np.cross(e1, e2).astype(np.float64)
    fn /= np.maximum(np.linalg.norm(fn, axis=1, keepdims=True), 1e-20)
    N = (fn[:, None, :] + rng.normal(0, 0.2, (n_tris, 3, 3))).astype(f32)
    T = rng.uniform(0, 1, (n_tris, 3, 2)).astype(f32)
    faces = np.zeros(n_tris, dtype=FACE_DTYPE)
    faces["v"], faces["vn"], faces["vt"] = V, N, T
    faces["smooth"] = 1
    faces["material"] = rng.integers(0, 16, n_tris)
    faces["texture"] = np.where(rng.random(n_tris) < 0.5, rng.integers(0, 4, n_tris), -1)
    spheres = np.zeros(n_spheres, dtype=SPHERE_DTYPE)
    spheres["center"] = rng.uniform(-50, 50, (n_spheres, 3)).astype(f32)
    spheres["radius"] = rng.uniform(0.5, 2.0, n_spheres).astype(f32)
    spheres["material"] = rng.integers(0, 16, n_spheres)
    spheres["texture"] = np.where(rng.random(n_spheres) < 0.5, rng.integers(0, 4, n_spheres), -1)
    # ids: spheres first then faces, like a scene file listing the spheres first
    spheres["id"] = np.arange(1, n_spheres + 1)
    faces["id"] = np.arange(n_spheres + 1, n_spheres + n_tris + 1)
    mats = np.zeros(16, dtype=MATERIAL_DTYPE)
    for k in range(16):
        od = rng.uniform(0.2, 1.0, 3); os_ = rng.uniform(0.5, 1.0, 3)
        if k < 4:
            mats[k] = material(od, os_, 0.1, 0.4, rng.uniform(0.2, 0.6), rng.uniform(20, 80),
                               (0.2, 0.5)[k % 2], (1.33, 1.5)[k // 2])
        else:
            mats[k] = material(od, os_, 0.15, 0.6, rng.uniform(0.0, 0.8) if k % 3 else 0.0, rng.uniform(5, 80))
    lights = np.zeros(n_lights, dtype=LIGHT_DTYPE)
    for k in range(n_lights):
        col = rng.uniform(0.15, 0.3, 3)
        if k < n_lights // 2:
            d = rng.normal(0, 1, 3); d[1] = -abs(d[1]) - 0.5
            d = (d / np.linalg.norm(d)).astype(f32)
            # 7 significant digits, like a scene file would carry (unit length matters)
            d = np.array([float(f"{x:.7g}") for x in d], dtype=f32)
            lights[k] = light(d, 0, col)
        else:
            lights[k] = light(rng.uniform(-80, 80, 3), 1, col)
    kinds = ["earth", "wood", "checker", "grass"]
    textures = [make_texture(kinds[k], tex_size, tex_size, k + 1) for k in range(4)]
    return Scene(width, height, eye=(0, 0, 140), viewdir=(0, 0, -1), updir=(0, 1, 0), hfov_deg=50,
                 bkg=(0.2, 0.3, 0.5), bkg_eta=1.0, materials=mats, spheres=spheres, faces=faces, lights=lights,
                 textures=textures)


# ---- branch-coverage scene (glass box + nested spheres + both light kinds) ------------------
def coverage_scene(width: int = 160, height: int = 120, seed: int = 3) -> Scene:
    """Small scene exercising what the example scenes do not combine: transparent
    FACES (the `goto SKIP_TRANS` rule), a glass sphere inside the box, textured
    smooth and flat triangles, a non-unit directional light and a point light,
    partial-opacity shadows, bkgcolor without eta."""
    rng = np.random.default_rng(seed)
    mats = np.array([
        material((0.9, 0.9, 0.9), (1, 1, 1), 0.2, 0.6, 0.0, 10),                 # 0 matte floor
        material((0.2, 0.6, 0.9), (1, 1, 1), 0.1, 0.3, 0.5, 40, 0.25, 1.33),     # 1 glass box
        material((1.0, 1.0, 1.0), (1, 1, 1), 0.1, 0.2, 0.6, 60, 0.2, 1.5),       # 2 glass sphere
        material((0.8, 0.3, 0.2), (1, 1, 1), 0.2, 0.6, 0.4, 20),                 # 3 opaque shiny
        material((0.7, 0.7, 0.2), (0.5, 0.5, 0.5), 0.3, 0.5, 0.3, 8, 0.6, 1.1),  # 4 half transparent
    ], dtype=MATERIAL_DTYPE)
    faces = []

    def quad(p0, p1, p2, p3, mat, tex=-1, smooth=0):
        uv = [(0, 0), (1, 0), (1, 1), (0, 1)]
        for a, b, c in ((0, 1, 2), (0, 2, 3)):
            f = np.zeros(1, dtype=FACE_DTYPE)[0]
            P = [p0, p1, p2, p3]
            f["v"] = [P[a], P[b], P[c]]
            f["vt"] = [uv[a], uv[b], uv[c]]
            nrm = np.cross(np.subtract(p1, p0), np.subtract(p3, p0)).astype(np.float64)
            f["vn"] = [nrm + rng.normal(0, 0.15, 3) for _ in range(3)]
            f["smooth"], f["material"], f["texture"] = smooth, mat, tex
            faces.append(f)

    quad((-8, -1, 2), (8, -1, 2), (8, -1, 22), (-8, -1, 22), 0, tex=0, smooth=0)     # textured floor
    lo, hi = np.array([-1.5, -0.9, 8.0]), np.array([1.5, 2.0, 11.0])                  # glass box
    x0, y0, z0 = lo; x1, y1, z1 = hi
    quad((x0, y0, z0), (x1, y0, z0), (x1, y1, z0), (x0, y1, z0), 1)
    quad((x1, y0, z1), (x0, y0, z1), (x0, y1, z1), (x1, y1, z1), 1)
    quad((x0, y0, z1), (x0, y0, z0), (x0, y1, z0), (x0, y1, z1), 1)
    quad((x1, y0, z0), (x1, y0, z1), (x1, y1, z1), (x1, y1, z0), 1)
    quad((x0, y1, z0), (x1, y1, z0), (x1, y1, z1), (x0, y1, z1), 1)
    quad((x0, y0, z1), (x1, y0, z1), (x1, y0, z0), (x0, y0, z0), 1)
    quad((3, -1, 12), (6, -1, 12), (6, 3, 13), (3, 3, 13), 3, tex=1, smooth=1)       # textured smooth panel
    faces = np.array(faces, dtype=FACE_DTYPE)
    spheres = np.zeros(5, dtype=SPHERE_DTYPE)
    spheres["center"] = [(0, 0.4, 9.5), (-3.5, 0.5, 9), (3.0, 0.2, 7.0), (0.0, 3.2, 10.0), (-3.5, 0.5, 9)]
    spheres["radius"] = [0.8, 1.4, 0.9, 0.7, 0.6]
    spheres["material"] = [2, 2, 3, 4, 3]          # sphere 4 is nested inside sphere 1
    spheres["texture"] = [-1, -1, 1, -1, -1]
    n_f = len(faces)
    faces["id"] = np.arange(1, n_f + 1)
    spheres["id"] = np.arange(n_f + 1, n_f + 6)
    lights = np.array([light((-1, -2, 1.5), 0, (0.7, 0.7, 0.6)), light((4, 6, 3), 1, (0.5, 0.5, 0.6))], dtype=LIGHT_DTYPE)
    return Scene(width, height, eye=(0, 2, -2), viewdir=(0, -0.15, 1), updir=(0, 1, 0), hfov_deg=60,
                 bkg=(0.3, 0.5, 0.8), bkg_eta=0.0, materials=mats, spheres=spheres, faces=faces, lights=lights,
                 textures=[make_texture("checker", 64, 64, 11), make_texture("earth", 128, 64, 12)])


def random_soup(n_tris: int, n_spheres: int, seed: int, width: int = 64, height: int = 48, extent: float = 10.0,
                n_lights: int = 2, transparent: bool = True) -> Scene:
    """Small random scenes for brute-force == LBVH == oracle tests."""
    s = stress_scene(n_tris, n_spheres, width, height, seed, n_lights=n_lights, tex_size=32)
    sc = extent / 50.0
    s.faces["v"] = (s.faces["v"] * sc).astype(np.float32)
    # triangles scaled with the scene but kept visible: grow them relative to the extent
    c = s.faces["v"].mean(axis=1, keepdims=True)
    s.faces["v"] = (c + (s.faces["v"] - c) * (4.0 if n_tris < 5000 else 1.5)).astype(np.float32)
    s.spheres["center"] = (s.spheres["center"] * sc).astype(np.float32)
    s.spheres["radius"] = (s.spheres["radius"] * sc * 1.5).astype(np.float32)
    s.lights["xyz"][s.lights["w"] != 0] *= sc
    s.eye = np.array([0, 0, 2.8 * extent], dtype=np.float32)
    if not transparent:
        s.materials["opacity"] = 1.0
    s._invalidate()
    return s
