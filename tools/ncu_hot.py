#!/usr/bin/env python
"""Top SASS instructions of one kernel in an .ncu-rep by stall samples, with the stall mix.
usage: ncu_hot.py report.ncu-rep kernel-substring [top-n]"""
import csv, io, subprocess, sys

rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
blocks, cur = [], None
for line in out.splitlines():
    if line.startswith('"Kernel Name"'):
        cur = {"name": next(csv.reader([line]))[1], "lines": []}
        blocks.append(cur)
    elif cur is not None:
        cur["lines"].append(line)
for b in blocks:
    if pat not in b["name"]:
        continue
    rows = list(csv.DictReader(io.StringIO("\n".join(b["lines"]))))
    tot = sum(int(r["# Samples"] or 0) for r in rows)
    inst = sum(int(r["Instructions Executed"] or 0) for r in rows)
    thr = sum(int(r["Thread Instructions Executed"] or 0) for r in rows)
    print(f"== {b['name']}: {tot} samples, {inst} warp instructions, {thr / max(inst, 1):.1f} threads/instr")
    stall_cols = [c for c in rows[0] if c.startswith("stall_") and "Not Issued" not in c]
    mix = {c: sum(int(r[c] or 0) for r in rows) for c in stall_cols}
    print("   stall mix:", ", ".join(f"{k[6:]} {v * 100 // max(tot, 1)}%" for k, v in sorted(mix.items(), key=lambda kv: -kv[1])[:8]))
    rows_i = list(enumerate(rows))
    for i, r in sorted(rows_i, key=lambda ir: -int(ir[1]["# Samples"] or 0))[:top]:
        s = int(r["# Samples"] or 0)
        why = sorted(((int(r[c] or 0), c[6:]) for c in stall_cols), reverse=True)[:2]
        print(f"{i:5d} {s * 100.0 / max(tot, 1):5.1f}%  thr {r['Avg. Threads Executed']:>5}  {r['Source'].strip():60s} {why[0][1]}:{why[0][0]} {why[1][1]}:{why[1][0]}")
    break
