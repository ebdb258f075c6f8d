#!/usr/bin/env python
"""Markdown tables for BASELINE.md / DESIGN.md from the bench lines committed under profiles/."""
import json, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
def load(name):
    try:
        return json.loads(open(os.path.join(ROOT, "profiles", name)).read().strip().splitlines()[-1])
    except OSError:
        return None
rows = [("C1", "four_spheres_512"), ("C2", "mesh_earth_1080p"), ("C3", "test7_1080p"), ("C4", "test7_4k"), ("C4", "house_4k"),
        ("C5", "stress_1M_4k"), ("C5", "stress_1M_8k")]
print("| cfg | workload | size | prims | lights | ms / frame (1 GPU) | Mrays/s (reference-equivalent) | Mrays/s walked | roofline `frac` (bound) | binding unit | end to end ms | CPU reference, 16 host cores | frame_check |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
for cfg, w in rows:
    d = load(f"r2_bench_1gpu_{w}.json")
    c = d["config"]; r = d["roofline"]; b = r.get("binding", {}); cb = d.get("cpu_baseline", {}); e = d.get("e2e", {})
    bind = f"issue {b.get('issue_active_pct') or 0:.0f} % x {b.get('lanes_per_inst') or 0:.1f}/32 lanes" if b.get("issue_active_pct") else "issue (see DESIGN 5.3)"
    cpu = cb.get("value", 0) * 1e6
    cpu_s = f"{cpu:,.0f} rays/s" if cpu < 1e5 else f"{cpu / 1e6:.1f} Mrays/s"
    print(f"| {cfg} | `{w}` | {c['width']}x{c['height']} | {c['triangles']:,} tri + {c['spheres']:,} sph | {c['lights']} | {d['ms_per_step']:.4g} | {d['value']:,.0f} | {d['rays_walked_per_s'] / 1e6:,.0f} | {r['frac']:.3f} ({r['bound']}) | {bind} | {e.get('ms_per_step', 0):.2f} | {cpu_s} | {d['frame_check']['result']} |")
print()
print("| workload | N = 1 | N = 2 | N = 4 | N = 8 | efficiency at 8 | end to end at 8 |")
print("|---|---|---|---|---|---|---|")
for w in ("stress_1M_4k", "stress_1M_8k", "test7_4k", "house_4k"):
    cells = []
    base = None
    for n in (1, 2, 4, 8):
        d = load(f"r2_scale_{w}_{n}gpu.json") or (load(f"r2_bench_1gpu_{w}.json") if n == 1 else None)
        if d is None:
            cells.append("—"); continue
        if n == 1: base = d["ms_per_step"]
        cells.append(f"{d['ms_per_step']:.4g} ms" + ("" if n == 1 else f" ({base / n / d['ms_per_step'] * 100:.1f} %)"))
        last = d
    eff = base / 8 / last["ms_per_step"] * 100
    e2e = (last.get("e2e") or {}).get("ms_per_step")
    print(f"| `{w}` | " + " | ".join(cells) + f" | {eff:.1f} % | {e2e:.2f} ms |" if e2e else f"| `{w}` | " + " | ".join(cells) + f" | {eff:.1f} % | — |")
