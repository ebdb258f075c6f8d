#!/usr/bin/env python
"""Developer script: render example / generated scenes on the GPU, compare each
with the CPU oracle and print a parity + timing table.  Not part of the product."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from simple_raytracer_b200 import abi, scenegen  # noqa: E402
from simple_raytracer_b200.renderer import DeviceScene  # noqa: E402
from tests import common, oracle_lib as O  # noqa: E402


def check(name, scene, accel=abi.RT_ACCEL_AUTO, depth=4, reps=3):
    t0 = time.time()
    raw, ids, ost = O.oracle_render(scene, max_depth=depth)
    t_cpu = time.time() - t0
    with DeviceScene(scene) as ds:
        st = ds.render(max_depth=depth, accel=accel, instrument=True)
        rgb, flags = ds.download()
        pid = ds.download_primary_ids()
        best = st.render_ms
        for _ in range(reps):
            best = min(best, ds.render(max_depth=depth, accel=accel).render_ms)
    cmp = common.compare_frames(raw, rgb, flags)
    id_mismatch = int((pid.reshape(-1) != ids).sum())
    rays = st.rays_primary + st.rays_shadow + st.rays_transmit + st.rays_reflect
    out = dict(scene=name, size=f"{scene.width}x{scene.height}", prims=scene.n_prims,
               accel="lbvh" if st.accel_used == abi.RT_ACCEL_LBVH else "brute", gpu_ms=round(best, 3),
               cpu_oracle_s=round(t_cpu, 3), rays=rays, rays_oracle=ost.rays_total, nodes=st.nodes_total,
               nodes_oracle=ost.nodes_total, mrays_s=round(rays / best / 1e3, 1), id_mismatch=id_mismatch,
               regrows=st.queue_regrows, **cmp)
    print(json.dumps(out), flush=True)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--set", default="examples")
    args = ap.parse_args()
    if args.set in ("examples", "all"):
        for name in common.ALL_SCENES:
            sc = common.example_scene(name)
            if sc.n_pixels > 600 * 600:
                sc = sc.with_size(540, 540)
            check(name, sc)
            if sc.n_prims > 0:
                check(name + " [lbvh]", sc, accel=abi.RT_ACCEL_LBVH)
    if args.set in ("generated", "all"):
        check("coverage", scenegen.coverage_scene(320, 240))
        check("coverage [lbvh]", scenegen.coverage_scene(320, 240), accel=abi.RT_ACCEL_LBVH)
        for n, m, seed in ((300, 20, 1), (2000, 50, 2)):
            sc = scenegen.random_soup(n, m, seed, 128, 96)
            check(f"soup{n} [brute]", sc, accel=abi.RT_ACCEL_BRUTE)
            check(f"soup{n} [lbvh]", sc, accel=abi.RT_ACCEL_LBVH)
        check("mesh_earth 64x32 @192x108", scenegen.mesh_earth(64, 32, 192, 108, tex_scale=1))
    if args.set in ("big", "all"):
        for (w, h) in ((1920, 1080), (3840, 2160)):
            sc = common.example_scene("reflection_transparency_tests/test7").with_size(w, h)
            with DeviceScene(sc) as ds:
                st = ds.render()
                ms = min(ds.render().render_ms for _ in range(5))
                print(json.dumps(dict(scene="test7", size=f"{w}x{h}", gpu_ms=round(ms, 3), nodes=st.nodes_total,
                                      regrows=st.queue_regrows)), flush=True)
        t0 = time.time()
        sc = scenegen.stress_scene(1_000_000, 1000, 1920, 1080)
        print("stress gen s", round(time.time() - t0, 2), flush=True)
        with DeviceScene(sc) as ds:
            t0 = time.time()
            st = ds.render()
            print("first render s", round(time.time() - t0, 3), st.as_dict(), flush=True)
            ms = min(ds.render().render_ms for _ in range(3))
            rays = st.rays_primary + st.rays_shadow + st.rays_transmit + st.rays_reflect
            print(json.dumps(dict(scene="stress_1M", size="1920x1080", gpu_ms=round(ms, 3), rays=rays,
                                  mrays_s=round(rays / ms / 1e3, 1))), flush=True)


if __name__ == "__main__":
    main()
