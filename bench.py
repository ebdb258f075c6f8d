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
from simple_raytracer_b200.multigpu import DistributedFrame  # noqa: E402
from simple_raytracer_b200.renderer import DeviceScene, make_params  # noqa: E402
from simple_raytracer_b200.scene import Scene  # noqa: E402

METRIC = "Mrays/s (primary+secondary)"
TILE_ROWS = 10          # 2160 / 10 = 216 tiles and 4320 / 10 = 432: both multiples of 8
L2_BYTES = 126 << 20


# ---- workloads (BASELINE.json configs) ------------------------------------------
def _example(name, w, h, tex_scale=1):
    from tests import common
    return common.example_scene(name, tex_scale).with_size(w, h)


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
    # configs[3] (ii): the showcase scene with its four textures (seeded 1024-wide stand-ins)
    "house_4k": lambda: _example("showcases/house", 3840, 2160, tex_scale=8),
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
_CPU_SCENE = None      # inherited by the forked workers (a Scene caches ctypes pointers: not picklable)


def _cpu_worker(args):
    pixels, kind = args
    scene = _CPU_SCENE
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
    if kind == "reference":
        # the forked workers trace through oracle/_ref/libref.so; one pixel in this process as well, so that
        # the library is mapped here too (the driver records the libraries of the bench process itself)
        O.load_ref()
        O.ref_render_pixels(scene, cpu_sample_pixels(scene, 1))
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
    # interleave the sample over the workers so that every process sees the same mix of pixels
    parts = [pixels[k::procs] for k in range(procs) if len(pixels[k::procs])]
    global _CPU_SCENE
    _CPU_SCENE = scene
    t0 = time.perf_counter()
    if len(parts) == 1:
        secs = [_cpu_worker((parts[0], kind))]
    else:
        with mp.get_context("fork").Pool(len(parts)) as pool:
            secs = pool.map(_cpu_worker, [(p, kind) for p in parts])
    wall = time.perf_counter() - t0
    if st is None:
        _, _, st = O.oracle_render(scene, pixels=pixels)
    rays = st.rays_total
    busy = max(secs)                     # slowest worker's time inside the reference's hot functions
    return {"value": rays / busy / 1e6, "unit": "Mrays/s", "cores": len(parts), "kind": kind,
            "sample": f"{len(pixels)} pixels on a regular grid of the {scene.width}x{scene.height} frame "
                      f"({rays} rays, {busy:.1f} s in the slowest of {len(parts)} processes, {wall:.1f} s wall incl. scene load)",
            "rays": int(rays), "seconds": busy}


# ---- peaks -----------------------------------------------------------------------------------
def measured_peaks() -> dict:
    """HBM GB/s from the driver-written MEASURED_PEAKS.json, else the profiling guide's fallback."""
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            d = json.load(f)
        return {"hbm_gbs": float(d["hbm_gbs"]), "sm_max_mhz": float(d.get("sm_max_mhz", 1965.0)), "source": "measured"}
    except (OSError, KeyError, ValueError):
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "source": "fallback"}


def fp32_peak_tflops(sm_mhz: float, sms: int = 148) -> float:
    """FP32 FMA peak: SMs x 128 lanes x 2 flop x clock."""
    return sms * 128 * 2 * sm_mhz * 1e6 / 1e12


# ---- pinned host copies of the scene arrays (the e2e leg uploads from pinned memory) ---------------
def pin_scene(scene: Scene):
    import torch
    keep = []

    def pin(a: np.ndarray) -> np.ndarray:
        if a.nbytes == 0:
            return a
        t = torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1)).pin_memory()
        keep.append(t)
        return t.numpy().view(a.dtype).reshape(a.shape)

    sc = Scene(scene.width, scene.height, scene.eye, scene.viewdir, scene.updir, scene.hfov_deg, scene.bkg,
               scene.bkg_eta, pin(scene.materials), pin(scene.spheres), pin(scene.faces), pin(scene.lights),
               [pin(t) for t in scene.textures])
    sc._pinned = keep
    return sc


# SURVEY 8(d) "Algorithmic work per unit": 32 B per node visited, 48 B per primitive tested, 64 B per ray /
# closest hit shaded, 4 B per texel.  (The trees here have 64-byte 4-wide and 96-byte 8-wide nodes and 80-byte
# primitive records, so the bytes that actually move per visit are larger; this line's figure is the
# implementation-independent one the survey fixed.)
B_NODE, B_PRIM, B_RAY, B_TEXEL = 32, 48, 64, 4


def algorithmic_work(scene: Scene, st, accel_used: int) -> dict:
    """algorithmic bytes (LBVH) or flops (brute force) of one frame, from the
    instrumented frame's device counters, split by kernel for the LBVH pipeline."""
    nodes = int(st.nodes_total)
    rays = int(st.rays_primary + st.rays_shadow + st.rays_transmit + st.rays_reflect)
    L = len(scene.lights)
    if accel_used == abi.RT_ACCEL_LBVH:
        sb, sp = int(st.shadow_boxes_tested), int(st.shadow_prim_tests)
        tb, tp = int(st.bvh_nodes_visited) - sb, int(st.prim_tests) - sp
        closest_rays = int(st.rays_primary + st.rays_transmit + st.rays_reflect)
        walked = int(st.shadow_rays_walked)
        per_kernel = {
            "bvh_trace_kernel": B_NODE * tb + B_PRIM * tp + B_RAY * closest_rays,
            "bvh_shadow_kernel": B_NODE * sb + B_PRIM * sp,
            "bvh_colour_kernel": B_RAY * nodes + B_TEXEL * int(st.texel_fetches),
            "bvh_spawn_kernel": B_RAY * (nodes + closest_rays - int(st.rays_primary)),      # hits read, child rays written
        }
        return {"bound": "hbm", "per_frame": sum(per_kernel.values()), "per_kernel": per_kernel, "unit": "GB/s",
                "detail": {"trace_nodes_visited": tb, "trace_prim_tests": tp, "shadow_nodes_visited": sb,
                           "shadow_prim_tests": sp, "closest_hit_rays": closest_rays, "shadow_rays_walked": walked,
                           "shadow_rays_reference": int(st.rays_shadow), "nodes_shaded": nodes,
                           "texels": int(st.texel_fetches), "shadow_chains_redone": int(st.shadow_chains_redone),
                           "shadow_rays_settled_by_occluder_map": int(st.shadow_map_hits),
                           "bvh_levels": [int(st.bvh_levels4), int(st.bvh_levels8)]}}
    fl = rays * (19 * len(scene.spheres) + 42 * len(scene.faces)) + nodes * (64 + 52 * L)
    return {"bound": "fp32", "per_frame": fl, "unit": "TFLOP/s", "detail": {"rays": rays, "nodes_shaded": nodes}}


def ncu_summary(workload: str) -> dict:
    """What the committed `ncu --set full` capture of this workload's dominant kernels says
    (profiles/r2_ncu_summary.json, written by tools/ncu_summary.py from the .ncu-rep): issue-slot use,
    lanes per instruction, DRAM bytes per launch.  A profiler cannot run inside the timed run, so these
    are the one part of the line that is not live; the file they come from is named beside them."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_ncu_summary.json")) as f:
            return json.load(f).get(workload, {})
    except (OSError, ValueError):
        return {}


def frame_sha256(t) -> str:
    import hashlib
    return hashlib.sha256(t.cpu().numpy().tobytes()).hexdigest()


# ---- our arm ---------------------------------------------------------------------------------
def run_ours(args) -> dict:
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"      # keep NCCL's banner out of stdout: rank 0 prints ONE JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    scene = WORKLOADS[args.workload]()
    W, H = scene.width, scene.height
    ds = DeviceScene(scene)
    # a real (non-default) stream: the C ABI reads a NULL stream as "the scene's own stream"
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sptr = stream.cuda_stream
    prm = make_params(rank=rank, world_size=world, tile_rows=TILE_ROWS)

    # one synchronous instrumented frame per rank: sizes the queues, counts rays and algorithmic work
    st_i = ds.render(rank=rank, world_size=world, tile_rows=TILE_ROWS, instrument=True)
    accel_used = st_i.accel_used
    rays_local = int(st_i.rays_primary + st_i.rays_shadow + st_i.rays_transmit + st_i.rays_reflect)
    work = algorithmic_work(scene, st_i, accel_used)
    # per-generation kernel times of a plain (un-instrumented) frame: roofline of the dominant kernel
    st_plain = ds.render(rank=rank, world_size=world, tile_rows=TILE_ROWS)
    st_p = ds.render(rank=rank, world_size=world, tile_rows=TILE_ROWS, profile=True)
    level_ms = [float(x) for x in list(st_p.level_ms)[:5]]
    # (profiled frames: one serial chain with an event after every kernel; shadow_ms = the occluder-map pre-pass
    # bvh_shadow_settle_kernel + bvh_shadow_kernel + the empty redo launch, shade_ms = bvh_colour_kernel)
    kernel_ms = {"bvh_trace_kernel": [float(x) for x in list(st_p.trace_ms)[:5]],
                 "bvh_spawn_kernel": [float(x) for x in list(st_p.spawn_ms)[:5]],
                 "bvh_shadow_kernel": [float(x) for x in list(st_p.shadow_ms)[:5]],
                 "bvh_colour_kernel": [float(x) for x in list(st_p.shade_ms)[:5]]}
    # (kernels of a plain frame as the timed loop runs it; the profiled frame is a serial chain per generation)
    launches_per_frame = int(st_plain.kernel_launches)

    df = DistributedFrame(ds, TILE_ROWS, rank, world)
    flush = torch.empty(2 * L2_BYTES, dtype=torch.uint8, device="cuda")

    def frame_step():
        df.step(sptr)       # rt_render_async -> one allgather (N > 1) -> rt_deinterleave_tiles

    for _ in range(args.warmup):
        flush.zero_()
        frame_step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for a, b in ev:
        flush.zero_()                 # L2 flush between timed iterations (outside the timed events)
        a.record(stream)
        frame_step()
        b.record(stream)
    barrier()
    total_ms = sum(a.elapsed_time(b) for a, b in ev)
    clocks = sampler.stop() if rank == 0 else None

    # ---- the frame the timed loop left behind: complete (no queue overflow on any rank), the same bytes on
    # every rank, and the bytes of an untimed single-GPU rt_render of the whole frame on rank 0
    complete = ds.async_frame_ok()
    my_hash = frame_sha256(df.frame)
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, (bool(complete), my_hash))
    else:
        gathered = [(bool(complete), my_hash)]
    frame_check = None
    if rank == 0:
        ds.render(rank=0, world_size=1, tile_rows=TILE_ROWS)
        rgb1, flg1 = ds.download()
        import hashlib
        single = hashlib.sha256(np.concatenate([rgb1, flg1[..., None]], axis=2).tobytes()).hexdigest()
        if not all(c for c, _ in gathered):
            frame_check = "INCOMPLETE: a wavefront queue overflowed on rank(s) " + str([k for k, (c, _) in enumerate(gathered) if not c])
        elif any(h != single for _, h in gathered):
            frame_check = "MISMATCH against the single-GPU frame on rank(s) " + str([k for k, (_, h) in enumerate(gathered) if h != single])
        else:
            frame_check = "byte-identical"
        frame_check = {"result": frame_check, "sha256": single, "ranks": world,
                       "what": "sha256 of every rank's assembled RGBA8 frame after the timed loop vs an untimed "
                               "single-GPU rt_render + rt_download_framebuffer on rank 0; rt_async_frame_status on every rank"}

    # rays the pipeline actually walks (the LBVH pipeline skips shadow rays whose result cannot matter)
    rays_walked_local = (work["detail"].get("closest_hit_rays", 0) + work["detail"].get("shadow_rays_walked", 0)) \
        if "closest_hit_rays" in work["detail"] else rays_local
    t = torch.tensor([total_ms, float(rays_local), float(rays_walked_local)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone(); dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        total_ms, rays_frame, rays_walked_frame = float(tmax[0]), int(tsum[1]), int(tsum[2])
    else:
        rays_frame, rays_walked_frame = rays_local, int(rays_walked_local)
    ms_per_step = total_ms / args.steps
    value = rays_frame / (ms_per_step * 1e-3) / 1e6

    # ---- e2e: the C ABI with host buffers, every step: upload (H2D from pinned memory, packing,
    # LBVH build) + render of this rank's tiles + download (D2H) of the rows this rank rendered
    e2e = None
    if not args.no_e2e:
        ds.close()
        pinned = pin_scene(scene)
        rgb_t = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()
        flg_t = torch.empty((H, W), dtype=torch.uint8).pin_memory()
        rgb, flg = rgb_t.numpy(), flg_t.numpy()
        n_e2e = max(1, min(args.steps, args.e2e_steps))

        split = {"upload_ms": 0.0, "build_plus_render_ms": 0.0, "download_ms": 0.0}

        def e2e_step():
            t0 = time.perf_counter()
            with DeviceScene(pinned) as d2:                 # rt_upload_scene: H2D of the description + SoA packing
                t1 = time.perf_counter()
                d2.render(prm)                              # rt_render: LBVH build (first frame of a scene) + the frame
                t2 = time.perf_counter()
                d2.download(rgb, flg)                       # rt_download_framebuffer: split planes + D2H
                t3 = time.perf_counter()
            split["upload_ms"] += (t1 - t0) * 1e3; split["build_plus_render_ms"] += (t2 - t1) * 1e3
            split["download_ms"] += (t3 - t2) * 1e3

        e2e_step()      # warm-up (first-touch of the pinned buffers, workspace cache)
        for k in split:
            split[k] = 0.0
        barrier()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            e2e_step()
        barrier()
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt[0])
        e2e = {"value": rays_frame * n_e2e / dt / 1e6, "unit": "Mrays/s", "ms_per_step": dt / n_e2e * 1e3,
               "steps": n_e2e, "h2d_bytes_per_step": pinned.h2d_bytes(),
               "d2h_bytes_per_step": int(H * W * 4 / world),
               "split_ms": {k: v / n_e2e for k, v in split.items()} | {"build_ms_est": max(0.0, split["build_plus_render_ms"] / n_e2e - ms_per_step)},
               "what": "rt_upload_scene(host RtSceneDesc, pinned) + rt_render + rt_download_framebuffer per step, "
                       "host wall clock, max over ranks"}
    else:
        ds.close()

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return {}

    peaks = measured_peaks()
    n_levels = 5
    ncu = ncu_summary(args.workload) if world == 1 else {}
    if work["bound"] == "hbm":
        # the dominant kernel of the LBVH pipeline: whichever tree-walk kernel takes longer
        dom = max(("bvh_trace_kernel", "bvh_shadow_kernel"), key=lambda k: sum(kernel_ms[k]))
        dom_ms = sum(kernel_ms[dom])
        achieved = work["per_kernel"][dom] / (dom_ms * 1e-3) / 1e9
        peak = peaks["hbm_gbs"]
        nk = ncu.get(dom, {})
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": nk.get("dram_bytes_per_launch"), "peak_source": peaks["source"], "kernel": dom,
                "algorithmic_bytes_per_launch": work["per_kernel"][dom] / n_levels,
                "per_unit_bytes": {"node_visit": B_NODE, "primitive_test": B_PRIM, "ray_or_shaded_hit": B_RAY, "texel": B_TEXEL,
                                   "source": "SURVEY.md 8(d)"},
                # what actually limits the kernel, from the committed ncu --set full capture of this workload
                "binding": {"unit": "issue slots at the SIMT efficiency of incoherent tree walks (not HBM)",
                            "issue_active_pct": nk.get("issue_active_pct"), "lanes_per_inst": nk.get("lanes_per_inst"),
                            "useful_lane_issue_frac": (nk["issue_active_pct"] / 100.0 * nk["lanes_per_inst"] / 32.0)
                            if nk.get("issue_active_pct") and nk.get("lanes_per_inst") else None,
                            "dram_gbs_measured": nk.get("dram_gbs"), "dram_pct_of_peak": nk.get("dram_pct"),
                            "l1_hit_pct": nk.get("l1_hit_pct"), "source": ncu.get("source")},
                "kernels": {k: {"ms_per_frame": sum(v), "ms_by_generation": v,
                                "algorithmic_bytes_per_frame": work["per_kernel"][k],
                                "achieved_gbs": work["per_kernel"][k] / (sum(v) * 1e-3) / 1e9 if sum(v) else None}
                            for k, v in kernel_ms.items()},
                "note": "frac = SURVEY 8(d) algorithmic bytes / kernel time / measured HBM copy bandwidth: the figure the survey "
                        "fixed for comparison.  Most of those bytes are served by L1/L2 (the upper tree is shared by all rays): "
                        "measured DRAM traffic is `traffic` per launch, a few percent of peak, and the kernel is bound by "
                        "instruction issue at ~18-20 of 32 lanes (`binding`), see DESIGN.md 6"}
    else:
        dom_ms = sum(level_ms)
        clk = (clocks or {}).get("sm_mhz") or peaks["sm_max_mhz"]
        peak = fp32_peak_tflops(clk)
        achieved = work["per_frame"] / (dom_ms * 1e-3) / 1e12
        nk = ncu.get("level_kernel", {})
        roof = {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "traffic": nk.get("dram_bytes_per_launch"),
                "peak_source": f"148 SM x 128 lanes x 2 x {clk:.0f} MHz (clock sampled during the run)",
                "kernel": "level_kernel (fused closest hit + shadow rays + shade + spawn)",
                "binding": {"unit": "exact-rounding FP32 divisions / FP64 conversions and libm (XU + FP64 pipes), launch latency on small frames",
                            "issue_active_pct": nk.get("issue_active_pct"), "lanes_per_inst": nk.get("lanes_per_inst"),
                            "xu_pipe_pct": nk.get("xu_pipe_pct"), "fp64_pipe_pct": nk.get("fp64_pipe_pct"),
                            "fma_pipe_pct": nk.get("fma_pipe_pct"), "source": ncu.get("source")}}
    roof.update({"launches_per_frame": n_levels, "ms_per_launch_avg": dom_ms / n_levels, "level_ms": level_ms,
                 "algorithmic_per_frame": work["per_frame"], "counters": work["detail"],
                 "frame_frac": work["per_frame"] / (float(st_p.render_ms) * 1e-3) / (1e9 if work["bound"] == "hbm" else 1e12) / peak
                 if st_p.render_ms else None,
                 "share_of_frame": dom_ms / float(st_p.render_ms) if st_p.render_ms else None,
                 "scope": "rank 0's shard" if world > 1 else "whole frame"})
    out = {"metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
           "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": {"workload": args.workload, "width": W, "height": H, "triangles": len(scene.faces),
                      "spheres": len(scene.spheres), "lights": len(scene.lights), "max_depth": 4},
           "config_note": {"accel": "lbvh" if accel_used == abi.RT_ACCEL_LBVH else "brute",
                           "rays_per_frame": rays_frame, "rays_walked_per_frame": rays_walked_frame,
                           "sharding": f"interleaved tiles of {TILE_ROWS} rows, "
                           f"{'NCCL allgather of RGBA8 tiles + ' if world > 1 else ''}device de-interleave",
                           "l2": "flushed between timed iterations (252 MiB memset, untimed)"},
           "rays_walked_per_s": rays_walked_frame / (ms_per_step * 1e-3),
           "value_note": "value counts the reference's TraceRay calls (primary + one shadow ray per light per shaded hit + "
                         "transmission + reflection); rays_walked_per_s counts the rays this pipeline had to walk",
           "frame_check": frame_check,
           "clocks": clocks, "gpu_launches": (launches_per_frame + 1) * args.steps, "roofline": roof}
    if e2e:
        out["e2e"] = e2e
    if world == 1 and not args.no_cpu:
        out["cpu_baseline"] = cpu_reference(scene, args.cpu_budget, os.cpu_count() or 1)
    if world > 1:
        dist.destroy_process_group()
    return out


# ---- reference arm ----------------------------------------------------------------------------
def run_reference(args) -> dict:
    """The reference's own CPU code for the path (oracle/_ref/libref.so, compiled from the
    unmodified sources; the plain-C port when that library is absent) on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return {}
    scene = WORKLOADS[args.workload]()
    cores = os.cpu_count() or 1
    steps = max(1, args.steps)
    # the whole --steps/--warmup run is bounded to ~args.cpu_budget * 4 seconds of wall time
    per_step = max(2.0, min(args.cpu_budget, 4.0 * args.cpu_budget / (steps + args.warmup)))
    for _ in range(min(args.warmup, 1)):
        cpu_reference(scene, min(per_step, 3.0), cores)
    runs = [cpu_reference(scene, per_step, cores) for _ in range(steps)]
    rays = sum(r["rays"] for r in runs)
    secs = sum(r["seconds"] for r in runs)
    value = rays / secs / 1e6
    base = runs[-1]
    return {"impl": "reference", "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": args.warmup, "ms_per_step": secs / steps * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": args.workload, "width": scene.width, "height": scene.height,
                       "triangles": len(scene.faces), "spheres": len(scene.spheres), "lights": len(scene.lights),
                       "max_depth": 4},
            "config_note": "the reference has no acceleration structure: brute force over all primitives per ray, "
                           "a bounded pixel sample of the frame per step",
            "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": base["cores"], "kind": base["kind"],
                             "sample": "each step: " + base["sample"]},
            "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for the cpu_baseline sample")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "ours":
        args.warmup = 3
    out = run_reference(args) if args.impl == "reference" else run_ours(args)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
