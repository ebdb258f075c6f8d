#!/usr/bin/env python
"""bench.py — throughput of the render hot path on 1..8 B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl ours|reference]

A "step" is one frame of the workload's scene.  `value` is whole-job Mrays/s with
the scene already resident in HBM (one ray = one TraceRay call of the reference:
primary + shadow + transmission + reflection); `e2e` is the same metric through
the C ABI with HOST buffers: rt_upload_scene (H2D of the scene description, SoA
packing, LBVH build) + rt_render + rt_download_framebuffer (D2H) every step.
For N > 1 (torchrun, one rank per GPU) every rank renders its interleaved row
tiles of the SAME frame, the 8-bit tiles are exchanged with one NCCL allgather and
de-interleaved on the device: strong scaling of one frame.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from simple_raytracer_b200 import abi, scenegen  # noqa: E402
from simple_raytracer_b200.renderer import DeviceScene, make_params  # noqa: E402
from simple_raytracer_b200.scene import Scene  # noqa: E402

METRIC = "Mrays/s (primary+secondary)"
TILE_ROWS = 10          # 2160 / 10 = 216 tiles and 4320 / 10 = 432: both multiples of 8
L2_BYTES = 126 << 20


# ---- workloads (BASELINE.json configs) ------------------------------------------
def _example(name, w, h):
    from tests import common
    return common.example_scene(name).with_size(w, h)


WORKLOADS = {
    # the north-star target: 4K, depth 5 generations, 1M textured triangles + 1k spheres, 8 lights, LBVH
    "stress_1M_4k": lambda: scenegen.stress_scene(1_000_000, 1000, 3840, 2160),
    # configs[4] at its full 8K size
    "stress_1M_8k": lambda: scenegen.stress_scene(1_000_000, 1000, 7680, 4320),
    # configs[1]: smooth textured mesh, directional + point light, 1920x1080
    "mesh_earth_1080p": lambda: scenegen.mesh_earth(512, 256, 1920, 1080),
    # configs[2] / configs[3]: the reference's own reflection / refraction demo scene
    "test7_1080p": lambda: _example("reflection_transparency_tests/test7", 1920, 1080),
    "test7_4k": lambda: _example("reflection_transparency_tests/test7", 3840, 2160),
    # configs[0]
    "four_spheres_512": lambda: _example("basic_geometry_tests/four_spheres", 512, 512),
    # small variant for quick checks of this script
    "stress_20k_480p": lambda: scenegen.stress_scene(20_000, 100, 640, 480, tex_size=64),
}
DEFAULT_WORKLOAD = "stress_1M_4k"


# ---- clocks ----------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[5 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---- CPU reference (the reference's own ShadeRay / TraceRay from oracle/_ref) -------------------
def _cpu_worker(args):
    scene, pixels, kind = args
    from tests import oracle_lib as O
    if kind == "reference":
        raw, ids, secs = O.ref_render_pixels(scene, pixels)
    else:
        t0 = time.perf_counter()
        O.oracle_render(scene, pixels=pixels)
        secs = time.perf_counter() - t0
    return secs


def cpu_sample_pixels(scene: Scene, n: int) -> np.ndarray:
    """n pixels on a regular grid over the whole frame (deterministic)."""
    gy = max(1, int(round((n * scene.height / scene.width) ** 0.5)))
    gx = max(1, n // gy)
    ys = ((np.arange(gy) + 0.5) * scene.height / gy).astype(np.int64)
    xs = ((np.arange(gx) + 0.5) * scene.width / gx).astype(np.int64)
    return (ys[:, None] * scene.width + xs[None, :]).reshape(-1)


def cpu_reference(scene: Scene, budget_s: float, procs: int) -> dict:
    """Times the reference's own code on a bounded pixel sample of `scene`.

    The reference is single-threaded and not re-entrant, so "all host
    threads" means `procs` independent processes, each tracing its share of the
    sample through its own copy of the reference (oracle/_ref/libref.so).  Rays are
    counted by the oracle port on the same pixels (bit-identical semantics)."""
    import multiprocessing as mp
    from tests import oracle_lib as O
    kind = "reference" if O.have_ref() else "port"
    # calibrate on a few pixels with the (much faster) port, then size the sample
    probe = cpu_sample_pixels(scene, 4)
    t0 = time.perf_counter()
    _, _, st = O.oracle_render(scene, pixels=probe)
    port_s = max(time.perf_counter() - t0, 1e-4)
    slowdown = 12.0 if kind == "reference" else 1.0       # reference / port cost ratio per ray.primitive (measured ~10x)
    per_pixel = port_s / len(probe) * slowdown
    n = int(max(procs, min(scene.n_pixels, budget_s * procs / per_pixel)))
    pixels = cpu_sample_pixels(scene, n)
    _, _, st = O.oracle_render(scene, pixels=pixels) if kind == "reference" else (None, None, None)
    parts = [p for p in np.array_split(pixels, procs) if len(p)]
    t0 = time.perf_counter()
    if len(parts) == 1:
        secs = [_cpu_worker((scene, parts[0], kind))]
    else:
        with mp.get_context("fork").Pool(len(parts)) as pool:
            secs = pool.map(_cpu_worker, [(scene, p, kind) for p in parts])
    wall = time.perf_counter() - t0
    if st is None:
        _, _, st = O.oracle_render(scene, pixels=pixels)
    rays = st.rays_total
    busy = max(secs)                     # slowest worker's time inside the reference's hot functions
    return {"value": rays / busy / 1e6, "unit": "Mrays/s", "cores": len(parts), "kind": kind,
            "sample": f"{len(pixels)} pixels on a regular grid of the {scene.width}x{scene.height} frame "
                      f"({rays} rays, {busy:.1f} s in the slowest of {len(parts)} processes, {wall:.1f} s wall incl. scene load)",
            "rays": int(rays), "seconds": busy}
