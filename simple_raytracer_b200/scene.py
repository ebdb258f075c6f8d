"""Host-side scene description: what the reference keeps in `Globals environment`
(src/definitions.h:304-311) plus main()'s camera locals (main.cpp:80-86), as flat
numpy arrays whose memory layout IS the C ABI layout of include/rt.h.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

from . import abi

MATERIAL_DTYPE = np.dtype([("od", "<f4", (3,)), ("os", "<f4", (3,)), ("ka", "<f4"), ("kd", "<f4"),
                           ("ks", "<f4"), ("n", "<f4"), ("opacity", "<f4"), ("eta", "<f4")])
SPHERE_DTYPE = np.dtype([("center", "<f4", (3,)), ("radius", "<f4"), ("material", "<i4"),
                         ("texture", "<i4"), ("id", "<i4")])
FACE_DTYPE = np.dtype([("v", "<f4", (3, 3)), ("vn", "<f4", (3, 3)), ("vt", "<f4", (3, 2)),
                       ("smooth", "<i4"), ("material", "<i4"), ("texture", "<i4"), ("id", "<i4")])
LIGHT_DTYPE = np.dtype([("xyz", "<f4", (3,)), ("w", "<f4"), ("rgb", "<f4", (3,))])

assert MATERIAL_DTYPE.itemsize == C.sizeof(abi.RtMaterial)
assert SPHERE_DTYPE.itemsize == C.sizeof(abi.RtSphere)
assert FACE_DTYPE.itemsize == C.sizeof(abi.RtFace)
assert LIGHT_DTYPE.itemsize == C.sizeof(abi.RtLight)


def _arr(a, dtype) -> np.ndarray:
    if a is None:
        return np.zeros(0, dtype=dtype)
    return np.ascontiguousarray(np.asarray(a, dtype=dtype))


class Scene:
    """A scene ready to be handed to rt_upload_scene (or to the test oracle)."""

    def __init__(self, width: int, height: int, eye, viewdir, updir, hfov_deg: float, bkg,
                 bkg_eta: float = 0.0, materials=None, spheres=None, faces=None, lights=None,
                 textures: Optional[Sequence[np.ndarray]] = None):
        self.width, self.height = int(width), int(height)
        self.eye = np.asarray(eye, dtype=np.float32).reshape(3)
        self.viewdir = np.asarray(viewdir, dtype=np.float32).reshape(3)
        self.updir = np.asarray(updir, dtype=np.float32).reshape(3)
        self.hfov_deg = float(np.float32(hfov_deg))
        self.bkg = np.asarray(bkg, dtype=np.float32).reshape(3)
        self.bkg_eta = float(np.float32(bkg_eta))
        self.materials = _arr(materials, MATERIAL_DTYPE)
        self.spheres = _arr(spheres, SPHERE_DTYPE)
        self.faces = _arr(faces, FACE_DTYPE)
        self.lights = _arr(lights, LIGHT_DTYPE)
        self.textures = [np.ascontiguousarray(np.asarray(t, dtype=np.uint8)) for t in (textures or [])]
        for t in self.textures:
            if t.ndim != 3 or t.shape[2] != 3:
                raise ValueError("textures must be [height][width][3] uint8")
        self._desc = None
        self._keep = None

    # -- counts ---------------------------------------------------------------
    @property
    def n_pixels(self) -> int:
        return self.width * self.height

    @property
    def n_prims(self) -> int:
        return len(self.spheres) + len(self.faces)

    def object_ids(self) -> np.ndarray:
        """Reference ids in TraceRay iteration order: faces first, then spheres."""
        return np.concatenate([self.faces["id"], self.spheres["id"]]).astype(np.int32)

    # -- C view -----------------------------------------------------------------
    def desc(self) -> abi.RtSceneDesc:
        if self._desc is not None:
            return self._desc
        d = abi.RtSceneDesc()
        d.width, d.height = self.width, self.height
        for k in range(3):
            d.eye[k], d.viewdir[k], d.updir[k], d.bkg[k] = self.eye[k], self.viewdir[k], self.updir[k], self.bkg[k]
        d.hfov_deg, d.bkg_eta = self.hfov_deg, self.bkg_eta
        tex = (abi.RtTexture * max(1, len(self.textures)))()
        for i, t in enumerate(self.textures):
            tex[i].height, tex[i].width = t.shape[0], t.shape[1]
            tex[i].rgb = t.ctypes.data_as(C.POINTER(C.c_uint8))
        d.n_materials = len(self.materials)
        d.materials = self.materials.ctypes.data_as(C.POINTER(abi.RtMaterial))
        d.n_textures = len(self.textures)
        d.textures = C.cast(tex, C.POINTER(abi.RtTexture))
        d.n_spheres = len(self.spheres)
        d.spheres = self.spheres.ctypes.data_as(C.POINTER(abi.RtSphere))
        d.n_faces = len(self.faces)
        d.faces = self.faces.ctypes.data_as(C.POINTER(abi.RtFace))
        d.n_lights = len(self.lights)
        d.lights = self.lights.ctypes.data_as(C.POINTER(abi.RtLight))
        self._desc, self._keep = d, tex
        return d

    def _invalidate(self):
        self._desc = None
        self._keep = None

    def h2d_bytes(self) -> int:
        """Bytes of scene description that rt_upload_scene reads from host memory."""
        return (self.materials.nbytes + self.spheres.nbytes + self.faces.nbytes + self.lights.nbytes +
                sum(t.nbytes for t in self.textures) + C.sizeof(abi.RtSceneDesc))

    # -- variants ---------------------------------------------------------------
    def copy(self, **overrides) -> "Scene":
        kw = dict(width=self.width, height=self.height, eye=self.eye.copy(), viewdir=self.viewdir.copy(),
                  updir=self.updir.copy(), hfov_deg=self.hfov_deg, bkg=self.bkg.copy(), bkg_eta=self.bkg_eta,
                  materials=self.materials.copy(), spheres=self.spheres.copy(), faces=self.faces.copy(),
                  lights=self.lights.copy(), textures=[t.copy() for t in self.textures])
        kw.update(overrides)
        return Scene(**kw)

    def with_size(self, width: int, height: int) -> "Scene":
        """Same scene with another `imsize`."""
        return self.copy(width=width, height=height)

    # -- parsing ----------------------------------------------------------------
    @classmethod
    def from_file(cls, path: str, texture_dir: Optional[str] = None) -> "Scene":
        """Parses a scene file with the host front-end (reference grammar, main.cpp:88-602)."""
        host = abi.load_host()
        out = C.POINTER(abi.RtSceneDesc)()
        rc = host.rt_host_parse_scene_file(os.fsencode(path), os.fsencode(texture_dir) if texture_dir else None,
                                           C.byref(out))
        abi.check(rc, f"rt_host_parse_scene_file({path})", host=True)
        try:
            d = out.contents

            def take(ptr, n, dtype):
                if n == 0:
                    return np.zeros(0, dtype=dtype)
                buf = C.string_at(ptr, n * dtype.itemsize)
                return np.frombuffer(buf, dtype=dtype).copy()

            textures = []
            for i in range(d.n_textures):
                t = d.textures[i]
                buf = C.string_at(t.rgb, t.width * t.height * 3)
                textures.append(np.frombuffer(buf, dtype=np.uint8).reshape(t.height, t.width, 3).copy())
            return cls(d.width, d.height, list(d.eye), list(d.viewdir), list(d.updir), d.hfov_deg, list(d.bkg),
                       d.bkg_eta, take(d.materials, d.n_materials, MATERIAL_DTYPE),
                       take(d.spheres, d.n_spheres, SPHERE_DTYPE), take(d.faces, d.n_faces, FACE_DTYPE),
                       take(d.lights, d.n_lights, LIGHT_DTYPE), textures)
        finally:
            host.rt_host_free_scene_desc(out)

    # -- writing (for feeding the same scene to the reference program) ------------
    def to_text(self, texture_paths: Optional[Sequence[str]] = None) -> str:
        """Scene-file text that parses back to this scene (one material per object
        run; floats printed with 9 significant digits so they round-trip)."""
        def f(x):
            return np.format_float_positional(np.float32(x), unique=True, trim="0")

        L = [f"eye {f(self.eye[0])} {f(self.eye[1])} {f(self.eye[2])}",
             f"viewdir {f(self.viewdir[0])} {f(self.viewdir[1])} {f(self.viewdir[2])}",
             f"updir {f(self.updir[0])} {f(self.updir[1])} {f(self.updir[2])}",
             f"hfov {f(self.hfov_deg)}", f"imsize {self.width} {self.height}"]
        bk = f"bkgcolor {f(self.bkg[0])} {f(self.bkg[1])} {f(self.bkg[2])}"
        if self.bkg_eta != 0.0:
            bk += f" {f(self.bkg_eta)}"
        L.append(bk)
        for l in self.lights:
            L.append("light " + " ".join(f(x) for x in [*l["xyz"], l["w"], *l["rgb"]]))
        objs = [(int(s["id"]), "s", i) for i, s in enumerate(self.spheres)] + \
               [(int(t["id"]), "f", i) for i, t in enumerate(self.faces)]
        objs.sort()
        cur = (None, None)
        nv = nvn = nvt = 0
        for _, kind, i in objs:
            o = self.spheres[i] if kind == "s" else self.faces[i]
            key = (int(o["material"]), int(o["texture"]))
            if key != cur:
                m = self.materials[key[0]]
                L.append("mtlcolor " + " ".join(f(x) for x in [*m["od"], *m["os"], m["ka"], m["kd"], m["ks"], m["n"],
                                                               m["opacity"], m["eta"]]))
                if key[1] >= 0:
                    if texture_paths is None:
                        raise ValueError("scene has textures: texture_paths required")
                    L.append(f"texture {texture_paths[key[1]]}")
                cur = key
            if kind == "s":
                L.append("sphere " + " ".join(f(x) for x in [*o["center"], o["radius"]]))
            else:
                for k in range(3):
                    L.append("v " + " ".join(f(x) for x in o["v"][k]))
                    L.append("vn " + " ".join(f(x) for x in o["vn"][k]))
                    L.append("vt " + " ".join(f(x) for x in o["vt"][k]))
                a = [nv + 1, nv + 2, nv + 3]
                nv += 3; nvn += 3; nvt += 3
                if o["smooth"]:
                    L.append("f " + " ".join(f"{x}/{x}/{x}" for x in a))
                else:
                    L.append("f " + " ".join(f"{x}/{x}" for x in a))
        return "\n".join(L) + "\n"
