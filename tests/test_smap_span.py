"""csrc/smap_span.h (the occluder-map builder's scanline spans, bvh.cu: build_shadow_maps) on the CPU: the header is
host/device code, compiled here with g++ and compared with the per-cell barycentric test on random triangles.
A disagreement may only be a cell whose centre lies on an edge within rounding (the map is a guess either way)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_row_spans_equal_the_per_cell_test(tmp_path):
    exe = str(tmp_path / "smap_span_check")
    subprocess.run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-I" + os.path.join(ROOT, "simple_raytracer_b200", "csrc"),
                    "-o", exe, os.path.join(ROOT, "tests", "native", "smap_span_check.cpp")], check=True)
    for seed in (1, 2, 3):
        out = subprocess.run([exe, str(seed), "30000"], check=True, capture_output=True, text=True).stdout.split()
        assert out[0] == "cells" and out[4] == "far_off", out
        both, span_only, test_only, far_off = int(out[1]), int(out[2]), int(out[3]), int(out[5])
        assert both > 500_000
        assert far_off == 0
        assert span_only + test_only <= both * 1e-4, out
