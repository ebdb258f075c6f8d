#!/usr/bin/env python
"""How many pixels of each example / generated scene are EXACTLY the reference's (the tests only demand
+-1 LSB on 99.9 %): GPU brute force and LBVH against the CPU oracle.  usage: exactness_report.py [out.txt]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from simple_raytracer_b200 import abi, scenegen
from simple_raytracer_b200.renderer import DeviceScene
from tests import common, oracle_lib as O

cases = [(n, common.golden_frame(n)[0]) for n in common.ALL_SCENES]
cases += [("Test1@full", common.golden_frame("reflection_transparency_tests/Test1@full")[0]),
          ("test7@full", common.golden_frame("reflection_transparency_tests/test7@full")[0]),
          ("coverage_scene", scenegen.coverage_scene(200, 150)), ("random_soup_300", scenegen.random_soup(300, 20, 1, 96, 72)),
          ("stress_8k", scenegen.stress_scene(8000, 100, 160, 120))]
lines = []
tot = exact = 0
for name, sc in cases:
    raw, ids, ost = O.oracle_render(sc)
    for accel, an in ((abi.RT_ACCEL_BRUTE, "brute"), (abi.RT_ACCEL_LBVH, "lbvh")):
        if sc.n_prims == 0 and accel == abi.RT_ACCEL_LBVH:
            continue
        with DeviceScene(sc) as ds:
            ds.render(accel=accel)
            rgb, flags = ds.download()
        c = common.compare_frames(raw, rgb, flags)
        tot += c["pixels"]; exact += c["exact"]
        lines.append(f"{name:44s} {an:6s} {sc.width}x{sc.height} exact {c['exact']}/{c['pixels']} max_err {c['max_err']} nan_ref {c['nan_ref']} nan_mismatch {c['nan_mismatch']}")
lines.append(f"TOTAL exact {exact}/{tot} = {exact / tot:.8f}")
out = "\n".join(lines)
print(out)
if len(sys.argv) > 1:
    open(sys.argv[1], "w").write(out + "\n")
