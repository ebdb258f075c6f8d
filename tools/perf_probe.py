#!/usr/bin/env python
"""Scratch perf probe: per-generation times and traversal counters of one workload."""
import argparse, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from simple_raytracer_b200 import abi, scenegen
from simple_raytracer_b200.renderer import DeviceScene

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--spheres", type=int, default=1000)
    ap.add_argument("--w", type=int, default=1920)
    ap.add_argument("--h", type=int, default=1080)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--depth", type=int, default=4)
    ap.add_argument("--no-instrument", action="store_true")
    ap.add_argument("--nop", action="store_true", help="no per-kernel profiling events (the frame then runs overlapped)")
    ap.add_argument("--world", type=int, default=1)
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--tile-rows", type=int, default=10)
    ap.add_argument("--subframes", type=int, default=0)
    a = ap.parse_args()
    sc = scenegen.stress_scene(a.tris, a.spheres, a.w, a.h)
    t0 = time.time()
    with DeviceScene(sc) as ds:
        up = time.time() - t0
        t0 = time.time()
        kw = dict(max_depth=a.depth, world_size=a.world, rank=a.rank, tile_rows=a.tile_rows)
        st = ds.render(**kw)
        first = time.time() - t0
        rays = st.rays_primary + st.rays_shadow + st.rays_transmit + st.rays_reflect
        best = None
        for _ in range(a.reps):
            s = ds.render(profile=not a.nop, subframes=a.subframes, **kw)
            if best is None or s.render_ms < best.render_ms:
                best = s
        out = dict(upload_s=round(up, 3), first_render_s=round(first, 3), ms=round(best.render_ms, 3), rays=int(rays),
                   mrays_s=round(rays / best.render_ms / 1e3, 1), nodes=list(best.nodes_per_level)[:a.depth + 1],
                   level_ms=[round(x, 3) for x in list(best.level_ms)[:a.depth + 1]], combine_ms=round(best.combine_ms, 3), trace_ms=[round(x, 3) for x in list(best.trace_ms)[:a.depth + 1]], spawn_ms=[round(x, 3) for x in list(best.spawn_ms)[:a.depth + 1]], shadow_ms=[round(x, 3) for x in list(best.shadow_ms)[:a.depth + 1]], shade_ms=[round(x, 3) for x in list(best.shade_ms)[:a.depth + 1]],
                   launches=best.kernel_launches, subframes=a.subframes)
        if not a.no_instrument:
            si = ds.render(instrument=True, **kw)
            out.update(boxes_per_ray=round(si.bvh_nodes_visited / rays, 1), prims_per_ray=round(si.prim_tests / rays, 2),
                       texels=int(si.texel_fetches), shadow_walked=int(si.shadow_rays_walked), shadow_ref=int(si.rays_shadow), instr_ms=round(si.render_ms, 2))
            for name in ("trace_span_ms", "trace_drain_ms", "trace_warp_drain_ms", "shadow_span_ms", "shadow_drain_ms", "shadow_warp_drain_ms",
                         "trace_drain_steps", "trace_drain_lanes", "trace_drain_steals", "shadow_drain_steps", "shadow_drain_lanes"):
                out[name] = [round(x, 3) for x in list(getattr(si, name))[:a.depth + 1]]
        print(json.dumps(out), flush=True)

if __name__ == "__main__":
    main()
